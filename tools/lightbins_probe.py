#!/usr/bin/env python3
"""Light bins of the shared-memory tracer: list-length histogram per light and, with the RS_LG_TIMING variant
(tools/build_variant.sh lgtime -DRS_LG_TIMING; RTB_LIB_PATH=build/variants/lib_lgtime.so), the cycles per coarse cell.
usage: lightbins_probe.py [view] [hours]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np  # noqa: E402
import torch  # noqa: E402

import raytracing_cuda_b200 as rt  # noqa: E402

view = int(sys.argv[1]) if len(sys.argv) > 1 else 5
hours = float(sys.argv[2]) if len(sys.argv) > 2 else 14.0
r = rt.Renderer(0)
r.set_sky(rt.synthetic_sky(512, 256))
s = rt.Scene()
f = s.configure(view=view, hours=hours, imgw=1920, imgh=1080, alias=True)
r.set_scene(s.objects, s.lights)
out = torch.empty(1920 * 1080, dtype=torch.int32, device="cuda")
for _ in range(3):
    r.render(f, out)
b = r.light_bins()
print("lights", [tuple(float(v) for v in l["pos"]) for l in s.lights])
for li in range(len(s.lights)):
    c = b[li, ..., 0].astype(np.int64)
    c[c == 0xFFFF] = 8
    print("light %d: bins by list length 0..7, overflow: %s   non-empty per face: %s"
          % (li, np.bincount(c.ravel(), minlength=9).tolist(), [(c[fa] > 0).sum() for fa in range(6)]))
if "lgtime" in os.environ.get("RTB_LIB_PATH", ""):
    t = b[3].reshape(-1).view(np.uint32).reshape(-1, 4)[: len(s.lights) * 96 * 8]
    cyc = t[:, 0].reshape(-1, 8)
    print("cycles per coarse cell (max over its warps): median %d, p90 %d, max %d" %
          (np.median(cyc.max(1)), np.percentile(cyc.max(1), 90), cyc.max()))
    worst = np.argsort(-cyc.max(1))[:8]
    for c in worst:
        print("  cell %3d (light %d face %d cell %2d): cycles %s  survivors %d  per warp %s" %
              (c, c // 96, (c % 96) // 16, c % 16, cyc[c].tolist(), t[c * 8, 1], t[c * 8:(c + 1) * 8, 2].tolist()))
