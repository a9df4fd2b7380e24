#!/usr/bin/env python3
"""Per-patch cost map of the persistent tracer (needs the -DRS_COSTMAP variant: tools/build_variant.sh cm -DRS_COSTMAP,
RTB_LIB_PATH=build/variants/lib_cm.so): ns per 8x4 patch, stored over the hit id of the patch's first pixel."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np, torch
import raytracing_cuda_b200 as rt
from raytracing_cuda_b200 import _lib
w, h = 1920, 1080
for view in (5, 6):
    r = rt.Renderer(0); r.set_sky(rt.synthetic_sky(2048, 1024))
    s = rt.Scene(); f = s.configure(view=view, hours=14.0, imgw=w, imgh=h, alias=True); r.set_scene(s.objects, s.lights)
    out = torch.empty(w * h, dtype=torch.int32, device="cuda"); ids = torch.zeros(w * h, dtype=torch.int32, device="cuda")
    for _ in range(3): r.render(f, out, hit_ids=ids)
    r.synchronize()
    cm = ids.cpu().numpy().reshape(h, w)[0::4, 0::8].astype(np.float64) / 1e3     # us per patch
    print("view %d: patches %s, mean %.2f us, p50 %.2f p95 %.2f p99 %.2f max %.2f" % (view, cm.shape, cm.mean(), np.median(cm),
          np.percentile(cm, 95), np.percentile(cm, 99), cm.max()))
    print(" by band of 27 patch rows (108 px): mean / max us")
    for b in range(0, cm.shape[0], 27):
        band = cm[b:b + 27]
        print("   rows %4d-%4d: mean %6.2f max %6.2f  share of total %.3f" % (b * 4, (b + 27) * 4 - 1, band.mean(), band.max(), band.sum() / cm.sum()))
    top = np.argsort(cm.ravel())[-12:][::-1]
    print(" slowest patches (x, y, us):", [(int(t % cm.shape[1]) * 8, int(t // cm.shape[1]) * 4, round(float(cm.ravel()[t]), 1)) for t in top])
    r.close()
