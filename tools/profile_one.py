#!/usr/bin/env python3
"""Render a few frames of one configuration (for ncu captures).  usage: profile_one.py [view] [hours] [w] [h] [frames] [trace_impl]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

import raytracing_cuda_b200 as rt  # noqa: E402
from raytracing_cuda_b200 import _lib, configs  # noqa: E402

a = sys.argv[1:]
view = int(a[0]) if len(a) > 0 else 5
hours = float(a[1]) if len(a) > 1 else 14.0
w = int(a[2]) if len(a) > 2 else 1920
h = int(a[3]) if len(a) > 3 else 1080
frames = int(a[4]) if len(a) > 4 else 6
impl = int(a[5]) if len(a) > 5 else 0
r = rt.Renderer(0)
r.set_sky(configs.sky_maps(os.path.join(ROOT, "oracle", "_ref", "backgrounds"))[0])     # the bench's sky maps
s = rt.Scene()
f = s.configure(view=view, hours=hours, imgw=w, imgh=h, alias=True)
r.set_scene(s.objects, s.lights)
r.set_option(_lib.RT_OPT_TRACE_IMPL, impl)
out = torch.empty(w * h, dtype=torch.int32, device="cuda")
r.set_option(_lib.RT_OPT_TIMING, 1)
for i in range(frames):
    r.render(f, out)
    r.synchronize()
    print(r.timings())
