#!/usr/bin/env python3
"""Where the time of a frame goes BETWEEN the kernels: %globaltimer stamps of pre-pass, tracer and FXAA over frames
rendered back to back on one context, as bench.py's device loop does (needs the -DRS_TIMELINE variant:
tools/build_variant.sh tl -DRS_TIMELINE; RTB_LIB_PATH=build/variants/lib_tl.so).

usage: frame_timeline.py [config2|config3] [frames<=8]
Prints, per frame, microseconds from the frame's first pre-pass CTA entry."""
import ctypes as C
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np  # noqa: E402
import torch  # noqa: E402

import raytracing_cuda_b200 as rt  # noqa: E402
from raytracing_cuda_b200 import _lib, configs  # noqa: E402

which = sys.argv[1] if len(sys.argv) > 1 else "config2"
n_frames = min(8, int(sys.argv[2])) if len(sys.argv) > 2 else 8
dbg = C.CDLL(_lib.LIB_PATH)
for fn in (dbg.rt_debug_timeline_small, dbg.rt_debug_timeline_fxaa):
    fn.argtypes = [C.c_void_p, C.c_int]
    fn.restype = C.c_int
sky, _ = configs.sky_maps(os.path.join(ROOT, "oracle", "_ref", "backgrounds"))
state = configs.config2()[1] if which == "config2" else configs.config3(sea_levels=(3.5,))[0]
_, objs, lights, f = state
r = rt.Renderer(0)
r.set_sky(sky)
r.set_scene(objs, lights)
out = torch.empty(f.imgw * f.imgh, dtype=torch.int32, device="cuda")
small = np.zeros((8, 8), dtype=np.uint64)
fx = np.zeros((8, 4), dtype=np.uint64)
for rep in range(3):
    for _ in range(10):
        r.render(f, out)
    torch.cuda.synchronize()
    assert dbg.rt_debug_timeline_small(small.ctypes.data, 1) == 0 and dbg.rt_debug_timeline_fxaa(fx.ctypes.data, 1) == 0
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n_frames):
        r.render(f, out)
    e1.record()
    torch.cuda.synchronize()
    assert dbg.rt_debug_timeline_small(small.ctypes.data, 0) == 0 and dbg.rt_debug_timeline_fxaa(fx.ctypes.data, 0) == 0
    print("%s %dx%d, %d frames back to back: %.1f us per frame by events" % (which, f.imgw, f.imgh, n_frames,
                                                                              1e3 * e0.elapsed_time(e1) / n_frames))
    s, x = small.astype(np.int64), fx.astype(np.int64)
    prev_end = None
    for i in range(n_frames):
        t0 = s[i, 0]
        us = lambda v: (v - t0) / 1e3
        line = ("frame %d: since previous FXAA end %6.1f | pre-pass done %5.1f | tracer CTA entry %5.1f, past wait %5.1f..%5.1f, "
                "last unit %6.1f, last exit %6.1f, late CTAs %d | FXAA entry %6.1f, past wait %6.1f, last exit %6.1f"
                % (i, float("nan") if prev_end is None else (t0 - prev_end) / 1e3, us(s[i, 1]), us(s[i, 2]), us(s[i, 3]),
                   us(s[i, 4]), us(s[i, 5]), us(s[i, 6]), s[i, 7], us(x[i, 0]), us(x[i, 1]), us(x[i, 2])))
        print(line)
        prev_end = x[i, 2]
