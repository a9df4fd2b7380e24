#!/usr/bin/env python3
"""Render a few frames of ONE shard of the built-in scene at 4K (for ncu launch lists).
usage: shard_one.py n_shards tile_w tile_h [frames]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import raytracing_cuda_b200 as rt
n, tw, th = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3])
frames = int(sys.argv[4]) if len(sys.argv) > 4 else 3
w, h = 3840, 2160
r = rt.Renderer(0); r.set_sky(rt.synthetic_sky(2048, 1024))
out = torch.empty(w * h, dtype=torch.int32, device="cuda")
s = rt.Scene(); f = s.configure(view=6, hours=14.0, imgw=w, imgh=h, alias=True); r.set_scene(s.objects, s.lights)
sh = rt.make_shard(n, n - 1, tw, th) if n > 1 else None
for _ in range(frames):
    r.render(f, out, shard=sh)
r.synchronize()
