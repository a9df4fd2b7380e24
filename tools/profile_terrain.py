#!/usr/bin/env python3
"""Render a few frames of config 4 (1M-triangle synthetic terrain, 3840x2160, camera looking down) for timing and
ncu captures.  usage: profile_terrain.py [frames] [w h] [cells]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

import raytracing_cuda_b200 as rt  # noqa: E402
from raytracing_cuda_b200 import _lib  # noqa: E402

a = sys.argv[1:]
frames = int(a[0]) if len(a) > 0 else 6
w, h = (int(a[1]), int(a[2])) if len(a) > 2 else (3840, 2160)
cells = int(a[3]) if len(a) > 3 else 708
r = rt.Renderer(0)
r.set_sky(rt.synthetic_sky(2048, 1024))
s = rt.Scene()
s.set_pose((324.4, 330.0, -84.0), 141.2, 35.0)
s.set_time(14.0)
s.set_antialiasing(True)
f = s.frame(w, h)
objs = rt.terrain_objects(cells)
r.set_option(_lib.RT_OPT_TIMING, 1)
r.set_option(_lib.RT_OPT_SHADOW_BINS, 2)                                  # light grid with the first frame
r.set_option(_lib.RT_OPT_WIDE_BVH, int(os.environ.get("RTB_WIDE", "0")))  # 1: the 4-wide view of the hierarchy
r.set_option(_lib.RT_OPT_TRACE_IMPL, int(os.environ.get("RTB_IMPL", "0")))
r.set_scene(objs, s.lights)
out = torch.empty(w * h, dtype=torch.int32, device="cuda")
r.set_option(_lib.RT_OPT_COUNTERS, 1)
r.render(f, out)
r.synchronize()
c = r.counters()
print("objects %d | per pixel: node tests %.1f tri %.2f | rays/px %.3f" % (len(objs), c["node_tests"] / (w * h),
      c["tri_tests"] / (w * h), rt.rays_per_frame(c) / (w * h)))
r.set_option(_lib.RT_OPT_COUNTERS, 0)
for i in range(frames):
    r.render(f, out)
    r.synchronize()
    t = r.timings()
    print("trace %.3f ms fxaa %.3f ms bvh %.3f ms | %.2f Grays/s" % (t["trace_ms"], t["fxaa_ms"], t["bvh_ms"],
          rt.rays_per_frame(c) / (t["trace_ms"] + t["fxaa_ms"]) / 1e6))
