#!/usr/bin/env python3
"""Per-warp timeline of the persistent tracer (needs the -DRS_TIMELINE variant: tools/build_variant.sh tl -DRS_TIMELINE,
RTB_LIB_PATH=build/variants/lib_tl.so).  The variant writes (start, last unit done, exit, patches) per warp into the
hit-id buffer AFTER the frame, so ids are garbage in this mode."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np, torch
import raytracing_cuda_b200 as rt
from raytracing_cuda_b200 import _lib
for (w, h, view) in ((1280, 720, 6), (1920, 1080, 5), (3840, 2160, 6)):
    r = rt.Renderer(0); r.set_option(_lib.RT_OPT_ORDER, int(os.environ.get('RTB_ORDER', '0'))); r.set_sky(rt.synthetic_sky(2048, 1024))
    s = rt.Scene(); f = s.configure(view=view, hours=14.0, imgw=w, imgh=h, alias=True); r.set_scene(s.objects, s.lights)
    out = torch.empty(w * h, dtype=torch.int32, device="cuda"); ids = torch.zeros(w * h, dtype=torch.int32, device="cuda")
    for _ in range(3): r.render(f, out, hit_ids=ids)
    r.synchronize()
    nw = 148 * 4 * 8
    t = ids.cpu().numpy().view(np.uint64)[:4 * nw].reshape(nw, 4).astype(np.float64)
    t = t[(t[:, 0] > 1e15) & (t[:, 1] > 1e15) & (t[:, 3] < 1e6) & (t[:, 0] < 1e19) & (t[:, 1] < 1e19) & (t[:, 2] < 1e19)]
    t = t[np.abs(t[:, 0] - np.median(t[:, 0])) < 1e9]; t = t[np.abs(t[:, 2] - np.median(t[:, 2])) < 1e9]; t = t[np.abs(t[:, 1] - np.median(t[:, 1])) < 1e9]      # drop entries overwritten by real hit ids
    nw = len(t)
    t0 = t[:, 0].min()
    st, last, ex, units = (t[:, 0] - t0) / 1e3, (t[:, 1] - t0) / 1e3, (t[:, 2] - t0) / 1e3, t[:, 3]
    print("%dx%d: warp start us min %.1f p50 %.1f max %.1f | last-unit-done p5 %.1f p50 %.1f p95 %.1f max %.1f | exit max %.1f | patches/warp min %d p50 %d max %d"
          % (w, h, st.min(), np.median(st), st.max(), np.percentile(last, 5), np.median(last), np.percentile(last, 95), last.max(), ex.max(), units.min(), np.median(units), units.max()))
    busy = (last - st)
    print("   per-patch us (busy/patches): p5 %.2f p50 %.2f p95 %.2f ; sum busy / (warps*span) = %.3f" % (
        np.percentile(busy / np.maximum(units, 1), 5), np.median(busy / np.maximum(units, 1)), np.percentile(busy / np.maximum(units, 1), 95), busy.sum() / (nw * ex.max())))
    tail = ex - last
    print("   after the last trace unit (FXAA phase) us: p5 %.1f p50 %.1f p95 %.1f max %.1f ; last-unit-done of the 20 latest warps: %s" % (
        np.percentile(tail, 5), np.median(tail), np.percentile(tail, 95), tail.max(), np.round(np.sort(last)[-20:], 1).tolist()))
    r.close()
