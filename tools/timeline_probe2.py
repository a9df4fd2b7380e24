#!/usr/bin/env python3
"""Per-warp timeline of the persistent shared-memory tracer on the 4K config-3 frame, whole and sharded (needs the
-DRS_TIMELINE variant: tools/build_variant.sh tl -DRS_TIMELINE; RTB_LIB_PATH=build/variants/lib_tl.so).  The variant
writes (kernel-entry time, last unit done, exit, units) per warp over the hit-id buffer after the frame."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np  # noqa: E402
import torch  # noqa: E402

import raytracing_cuda_b200 as rt  # noqa: E402
from raytracing_cuda_b200 import _lib, configs  # noqa: E402

w, h = 3840, 2160
r = rt.Renderer(0)
r.set_sky(configs.sky_maps(os.path.join(ROOT, "oracle", "_ref", "backgrounds"))[0])
_, objs, lights, f = configs.config3(sea_levels=(3.5,))[0]
r.set_scene(objs, lights)
out = torch.zeros(w * h, dtype=torch.int32, device="cuda")
ids = torch.zeros(w * h, dtype=torch.int32, device="cuda")
for n, tw, th in ((1, 256, 128), (2, 1920, 2160), (2, 3840, 1088), (2, 256, 128), (4, 256, 128), (8, 256, 128)):
    sh = rt.make_shard(n, 0, tw, th) if n > 1 else None
    for _ in range(4):
        r.render(f, out, hit_ids=ids, shard=sh)
    r.synchronize()
    r.set_option(_lib.RT_OPT_TIMING, 1)
    r.render(f, out, hit_ids=ids, shard=sh)
    r.synchronize()
    tm = r.timings()
    r.set_option(_lib.RT_OPT_TIMING, 0)
    nw = 148 * 4 * 8
    t = ids.cpu().numpy().view(np.uint64)[:4 * nw].reshape(nw, 4).astype(np.float64)
    ok = (t[:, 0] > 1e15) & (t[:, 0] < 1e19) & (t[:, 2] >= t[:, 0]) & (t[:, 2] - t[:, 0] < 1e9) & (t[:, 3] < 1e6)
    t = t[ok]
    t0 = t[:, 0].min()
    st, ex, units = (t[:, 0] - t0) / 1e3, (t[:, 2] - t0) / 1e3, t[:, 3]
    worked = units > 0
    last = np.where(worked, (t[:, 1] - t0) / 1e3, st)
    busy = last - st
    print("shards %d tile %dx%d: events: prepass+trace %.1f us | %d warps reported, %d got units | warp entry us p50 %.1f max %.1f | "
          "last unit done p5 %.1f p50 %.1f p95 %.1f max %.1f | units/warp p50 %d max %d | us per unit (busy/units) p50 %.2f | "
          "busy fraction of (entry..latest done) %.3f" % (
              n, tw, th, 1e3 * tm["trace_ms"], len(t), int(worked.sum()), np.median(st), st.max(), np.percentile(last[worked], 5),
              np.median(last[worked]), np.percentile(last[worked], 95), last.max(), np.median(units), units.max(),
              np.median(busy[worked] / units[worked]), busy.sum() / (len(t) * (last.max() - np.median(st)))), flush=True)
