#!/usr/bin/env python3
"""Kernel-time probe: trace / fxaa / build ms and test counts per variant, on the bench workload's frames.
usage: perf_probe.py [w h] ; prints one line per (view, time, builder)."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

import raytracing_cuda_b200 as rt  # noqa: E402
from raytracing_cuda_b200 import _lib  # noqa: E402

w, h = (int(sys.argv[1]), int(sys.argv[2])) if len(sys.argv) > 2 else (1920, 1080)
r = rt.Renderer(0)
r.set_sky(rt.synthetic_sky(2048, 1024))
s = rt.Scene()
r.set_option(_lib.RT_OPT_ORDER, int(os.environ.get("RTB_ORDER", "0")))
r.set_option(_lib.RT_OPT_SHADOW_BINS, int(os.environ.get("RTB_SHADOW_BINS", "0")))
r.set_option(_lib.RT_OPT_MIRROR_BINS, int(os.environ.get("RTB_MIRROR_BINS", "0")))
out = torch.empty(w * h, dtype=torch.int32, device="cuda")
for view in (5, 6):
    for tod in ("afternoon", "night"):
        for builder in ((1, 2) if os.environ.get("RTB_BOTH_BUILDERS") else (2,)):
            r.set_option(_lib.RT_OPT_BUILDER, builder)
            f = s.configure(view=view, hours=rt.TIME_PRESETS[tod], imgw=w, imgh=h, alias=True)
            r.set_scene(s.objects, s.lights)
            r.set_option(_lib.RT_OPT_COUNTERS, 1)
            r.render(f, out)
            c = r.counters()
            r.set_option(_lib.RT_OPT_COUNTERS, 0)
            r.set_option(_lib.RT_OPT_TIMING, 1)
            tr = fx = 0.0
            n = 20
            for _ in range(3):
                r.render(f, out)
            for _ in range(n):
                r.render(f, out)
                t = r.timings()
                tr += t["trace_ms"]
                fx += t["fxaa_ms"]
            r.set_option(_lib.RT_OPT_TIMING, 0)
            rays = rt.rays_per_frame(c)
            print("view %d %-9s builder %s: trace %.3f ms fxaa %.4f ms | %.2f Grays/s | per pixel: nodes %.1f tri %.2f sph %.2f"
                  % (view, tod, "karras" if builder == 1 else "ploc  ", tr / n, fx / n, rays / (tr / n + fx / n) / 1e6,
                     c["node_tests"] / (w * h), c["tri_tests"] / (w * h), c["sphere_tests"] / (w * h)), flush=True)
