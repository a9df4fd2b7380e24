#!/usr/bin/env python3
"""How does one rank's sharded render time scale with the shard count (logical shards on ONE GPU)?"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import raytracing_cuda_b200 as rt
from raytracing_cuda_b200 import _lib
w, h = 3840, 2160
r = rt.Renderer(0); r.set_sky(rt.synthetic_sky(2048, 1024))
out = torch.empty(w * h, dtype=torch.int32, device="cuda")
def t(fn, n=20):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(n): fn()
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) / n
for name in (sys.argv[1:] or ["builtin", "terrain"]):
    s = rt.Scene()
    if name == "builtin":
        f = s.configure(view=6, hours=14.0, imgw=w, imgh=h, alias=True); r.set_scene(s.objects, s.lights)
    else:
        s.set_pose((324.4, 330.0, -84.0), 141.2, 35.0); s.set_time(14.0); s.set_antialiasing(True); f = s.frame(w, h)
        r.set_scene(rt.terrain_objects(708), s.lights)
    for n, tw, th in ((1, 64, 32), (2, 64, 32), (2, 128, 64), (2, 256, 128), (4, 128, 64), (4, 256, 128), (8, 128, 64), (8, 256, 128), (8, 256, 128)):
        sh = rt.make_shard(n, n - 1, tw, th) if n > 1 else None
        r.set_option(_lib.RT_OPT_TIMING, 1)
        r.render(f, out, shard=sh); r.synchronize(); tm = r.timings()
        r.set_option(_lib.RT_OPT_TIMING, 0)
        ms = t(lambda: r.render(f, out, shard=sh))
        line = "%s shards=%d tile %dx%d: render %.3f ms (trace %.3f fxaa %.3f)" % (name, n, tw, th, ms, tm["trace_ms"], tm["fxaa_ms"])
        if sh is not None:
            packed = torch.zeros(r.tile_slots(sh, w, h) * tw * th, dtype=torch.int32, device="cuda")
            line += " pack %.3f unpack %.3f" % (t(lambda: r.pack_tiles(out, w, h, sh, packed)), t(lambda: r.unpack_tiles(packed, w, h, sh, out)))
        print(line, flush=True)
