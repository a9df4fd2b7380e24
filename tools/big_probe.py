#!/usr/bin/env python3
"""Config 4 (1,002,531-object terrain, 3840x2160, FXAA on) on one GPU: the plain one-thread-per-pixel tracer against
the persistent tracer with and without packet traversal of the primary rays.  One JSON line per variant."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

import raytracing_cuda_b200 as rt  # noqa: E402
from raytracing_cuda_b200 import _lib, configs  # noqa: E402

objs, lights, f = configs.config4()
r = rt.Renderer(0)
r.set_sky(rt.synthetic_sky(2048, 1024))
r.set_option(_lib.RT_OPT_SHADOW_BINS, 2)
r.set_scene(objs, lights)
out = torch.empty(f.imgw * f.imgh, dtype=torch.int32, device="cuda")
ref = None
for rnd in range(2):
    for name, impl, packet, sstack, wide in (("plain, binary LBVH", 1, 0, 0, 0), ("plain, 4-wide view", 1, 0, 0, 1),
                                             ("plain+packet, binary", 4, 0, 0, 0), ("plain+packet, 4-wide", 4, 0, 0, 1),
                                             ("persistent+packet", 3, 0, 0, 0),
                                             ("persistent, per-lane", 3, 1, 0, 0), ("persistent+packet, smem stack", 3, 0, 1, 0),
                                             ("persistent, per-lane, smem stack", 3, 1, 1, 0)):
        r.set_option(_lib.RT_OPT_WIDE_BVH, wide)
        r.set_option(_lib.RT_OPT_TRACE_IMPL, impl)
        r.set_option(_lib.RT_OPT_PACKET, packet)
        r.set_option(_lib.RT_OPT_BIG_STACK, sstack)
        r.set_option(_lib.RT_OPT_COUNTERS, 1)
        r.render(f, out)
        c = r.counters()
        r.set_option(_lib.RT_OPT_COUNTERS, 0)
        r.render(f, out)
        r.synchronize()
        if ref is None:
            ref = out.clone()
        same = bool(torch.equal(out, ref))
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
        px = f.imgw * f.imgh
        print(json.dumps({"variant": name, "round": rnd, "trace_ms": tr / n, "fxaa_ms": fx / n, "equals_plain": same,
                          "node_tests_per_px": c["node_tests"] / px, "tri_tests_per_px": c["tri_tests"] / px,
                          "device_errors": r.device_errors()}), flush=True)
