#!/usr/bin/env python3
"""Experiment: how much do tighter boxes for the big sloped triangles help?  Emulated at the scene level by
replacing each mountain triangle (objects 81..128) with its 4 (or 16) midpoint sub-triangles."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np, torch
import raytracing_cuda_b200 as rt
from raytracing_cuda_b200 import _lib

def split(objs, idxs, levels):
    out = []
    for i, o in enumerate(objs):
        if i not in idxs:
            out.append(o.copy()); continue
        tris = [(np.array(list(o["pos"]), np.float32), np.array(list(o["size"]), np.float32), np.array(list(o["third"]), np.float32))]
        for _ in range(levels):
            nt = []
            for a, b, c in tris:
                ab, bc, ca = (a + b) / 2, (b + c) / 2, (c + a) / 2
                nt += [(a, ab, ca), (ab, b, bc), (ca, bc, c), (ab, bc, ca)]
            tris = nt
        for a, b, c in tris:
            n = o.copy()
            n["pos"] = tuple(a); n["size"] = tuple(b); n["third"] = tuple(c)
            out.append(n)
    return np.array(out, dtype=objs.dtype)

w, h = 1920, 1080
r = rt.Renderer(0)
r.set_sky(rt.synthetic_sky(2048, 1024))
s = rt.Scene()
out = torch.empty(w * h, dtype=torch.int32, device="cuda")
for view in (5, 6):
    f = s.configure(view=view, hours=14.0, imgw=w, imgh=h, alias=True)
    base = s.objects.copy()
    for levels in (0, 1, 2):
        objs = split(base, set(range(81, 129)), levels) if levels else base
        for impl in (0, 1):
            r.set_option(_lib.RT_OPT_TRACE_IMPL, impl)
            r.set_scene(objs, s.lights)
            r.set_option(_lib.RT_OPT_COUNTERS, 1); r.render(f, out); c = r.counters(); r.set_option(_lib.RT_OPT_COUNTERS, 0)
            r.set_option(_lib.RT_OPT_TIMING, 1)
            for _ in range(3): r.render(f, out)
            tr = 0.0
            for _ in range(10):
                r.render(f, out); tr += r.timings()["trace_ms"]
            r.set_option(_lib.RT_OPT_TIMING, 0)
            print("view %d split-levels %d n=%d impl %d: trace %.3f ms | per pixel: nodes %.1f tri %.2f" % (
                view, levels, len(objs), impl, tr / 10, c["node_tests"] / (w * h), c["tri_tests"] / (w * h)), flush=True)
