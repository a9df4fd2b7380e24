#!/usr/bin/env python3
"""Per-pixel node-test map of the HBM tracer on the terrain scene (needs the -DRT_NODEMAP variant:
tools/build_variant.sh nm -DRT_NODEMAP, RTB_LIB_PATH=build/variants/lib_nm.so)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np, torch
import raytracing_cuda_b200 as rt
from raytracing_cuda_b200 import _lib
w, h = 3840, 2160
r = rt.Renderer(0); r.set_sky(rt.synthetic_sky(2048, 1024))
s = rt.Scene(); s.set_pose((324.4, 330.0, -84.0), 141.2, 35.0); s.set_time(14.0); s.set_antialiasing(True)
f = s.frame(w, h)
r.set_scene(rt.terrain_objects(708), s.lights)
out = torch.empty(w * h, dtype=torch.int32, device="cuda"); ids = torch.zeros(w * h, dtype=torch.int32, device="cuda")
r.set_option(_lib.RT_OPT_COUNTERS, 1)
r.render(f, out, hit_ids=ids); r.synchronize()
m = ids.cpu().numpy().reshape(h, w).astype(np.int64)
print("per pixel (node tests, or ns with the -DRT_TIMEMAP variant): mean %.1f p50 %d p99 %d p99.9 %d max %d" % (m.mean(), np.median(m), np.percentile(m, 99), np.percentile(m, 99.9), m.max()))
blk = m.reshape(h // 8, 8, w // 32, 32).max(axis=(1, 3))      # slowest lane per 32x8 block
print("per-block max: mean %.0f p99 %d max %d" % (blk.mean(), np.percentile(blk, 99), blk.max()))
top = np.argsort(m.ravel())[-10:][::-1]
print("heaviest pixels (x, y, node tests):", [(int(t % w), int(t // w), int(m.ravel()[t])) for t in top])
for b in range(0, h, 216):
    print("  rows %4d-%4d: mean %7.1f max %7d" % (b, b + 215, m[b:b + 216].mean(), m[b:b + 216].max()))
