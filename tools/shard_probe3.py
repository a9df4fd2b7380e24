#!/usr/bin/env python3
"""What does a shard of the 4K config-3 frame actually trace?  Counters (pixels, rays, node / leaf tests) and time of
shard 0 of N for several tile sizes, next to the whole frame.  One JSON line each."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

import raytracing_cuda_b200 as rt  # noqa: E402
from raytracing_cuda_b200 import _lib, configs  # noqa: E402

w, h = 3840, 2160
r = rt.Renderer(0)
r.set_sky(configs.sky_maps(os.path.join(ROOT, "oracle", "_ref", "backgrounds"))[0])
out = torch.zeros(w * h, dtype=torch.int32, device="cuda")
_, objs, lights, f = configs.config3(sea_levels=(3.5,))[0]
r.set_scene(objs, lights)
whole = None
for n, tw, th in ((1, 256, 128), (2, 256, 128), (2, 3840, 1088), (2, 1920, 2160), (2, 512, 256), (2, 128, 64), (2, 64, 32),
                  (4, 256, 128), (4, 512, 256), (8, 256, 128), (8, 128, 64)):
    sh = rt.make_shard(n, 0, tw, th) if n > 1 else None
    r.set_option(_lib.RT_OPT_COUNTERS, 1)
    r.render(f, out, shard=sh)
    c = r.counters()
    r.set_option(_lib.RT_OPT_COUNTERS, 0)
    r.set_option(_lib.RT_OPT_TIMING, 1)
    for _ in range(4):
        r.render(f, out, shard=sh)
    tr = 0.0
    for _ in range(20):
        r.render(f, out, shard=sh)
        tr += r.timings()["trace_ms"]
    r.set_option(_lib.RT_OPT_TIMING, 0)
    d = {"shards": n, "tile": "%dx%d" % (tw, th), "trace_us": 1e3 * tr / 20, "pixels": c["rays_primary"],
         "rays": rt.rays_per_frame(c), "node_tests": c["node_tests"], "leaf_tests": c["tri_tests"] + c["sphere_tests"]}
    if whole is None:
        whole = d
    for k in ("trace_us", "pixels", "rays", "node_tests", "leaf_tests"):
        d[k + "_share"] = round(d[k] / whole[k], 3)
    print(json.dumps(d), flush=True)
