#!/usr/bin/env python3
"""One rank's share of ONE 3840x2160 frame at 1 / 2 / 4 / 8 logical screen-tile shards on one GPU (256x128 tiles):
where does the time of a sharded frame go, and which scheduling options move it?  Config 3 (Mountains, sea 3.5) and
optionally config 4 (terrain).  One JSON line per (scene, shards, variant); times are the library's own events over
20 frames after warm-up (so that the order history of the shard is in place)."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

import raytracing_cuda_b200 as rt  # noqa: E402
from raytracing_cuda_b200 import _lib, configs  # noqa: E402

w, h = 3840, 2160
TW, TH = (int(v) for v in os.environ.get("RTB_TILE", "256x128").split("x"))
r = rt.Renderer(0)
r.set_sky(configs.sky_maps(os.path.join(ROOT, "oracle", "_ref", "backgrounds"))[0])
out = torch.zeros(w * h, dtype=torch.int32, device="cuda")
VARIANTS = {"default": {}, "compaction": {_lib.RT_OPT_COMPACTION: 1}, "bottom-up order": {_lib.RT_OPT_ORDER: 1}}
for name in (sys.argv[1:] or ["config3"]):
    if name == "config3":
        _, objs, lights, f = configs.config3(sea_levels=(3.5,))[0]
    else:
        objs, lights, f = configs.config4()
        r.set_option(_lib.RT_OPT_SHADOW_BINS, 2)
    r.set_scene(objs, lights)
    for n in (1, 2, 4, 8):
        for sid in sorted({0, n - 1}):
            sh = rt.make_shard(n, sid, TW, TH) if n > 1 else None
            for vname, opts in VARIANTS.items():
                if name != "config3" and vname != "default":
                    continue
                for k, v in opts.items():
                    r.set_option(k, v)
                r.set_option(_lib.RT_OPT_TIMING, 1)
                for _ in range(4):
                    r.render(f, out, shard=sh)
                tr = fx = 0.0
                reps = 20
                for _ in range(reps):
                    r.render(f, out, shard=sh)
                    t = r.timings()
                    tr += t["trace_ms"]
                    fx += t["fxaa_ms"]
                r.set_option(_lib.RT_OPT_TIMING, 0)
                for k in opts:
                    r.set_option(k, 0)
                print(json.dumps({"scene": name, "shards": n, "shard_id": sid, "tile": "%dx%d" % (TW, TH), "variant": vname,
                                  "prepass_plus_trace_us": 1e3 * tr / reps, "fxaa_us": 1e3 * fx / reps}), flush=True)
