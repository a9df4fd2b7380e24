#!/usr/bin/env python3
"""A/B probe of one rt_set_option key on the bench workload and on single frames, same process, same sky.

usage: ab_probe.py KEY V0,V1[,V2..] [reps] [config3]      e.g.  ab_probe.py 12 0,1     (RT_OPT_COMPACTION on / off)
       (a fourth argument "config3" runs the bench loop on the Mountains 4K sea-level frames instead of config 2)

  bench   bench.py's device loop: four contexts (Island 1080p, the four times of day), L2 flushed between steps, CUDA
          events per step -> Grays/s
  frame   one context, back-to-back frames of one view with the library's own trace / fxaa events
Prints one JSON line per setting."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

import raytracing_cuda_b200 as rt  # noqa: E402
from raytracing_cuda_b200 import _lib, configs  # noqa: E402

key = int(sys.argv[1])
values = [int(v) for v in sys.argv[2].split(",")]
reps = int(sys.argv[3]) if len(sys.argv) > 3 else 40
sky, label = configs.sky_maps(os.path.join(ROOT, "oracle", "_ref", "backgrounds"))
states = list(configs.config3()) if len(sys.argv) > 4 and sys.argv[4] == "config3" else configs.config2()
ctxs = []
for i, (_, objs, lights, fr) in enumerate(states):
    r = rt.Renderer(0)
    if i == 0:
        r.set_sky(sky)
    else:
        r.share_sky(ctxs[0][0])
    r.set_scene(objs, lights)
    ctxs.append((r, fr))
out = torch.empty(3840 * 2160, dtype=torch.int32, device="cuda")
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
rays = 0
for r, fr in ctxs:
    r.set_option(_lib.RT_OPT_COUNTERS, 1)
    r.render(fr, out)
    rays += rt.rays_per_frame(r.counters())
    r.set_option(_lib.RT_OPT_COUNTERS, 0)
stream = torch.cuda.current_stream()
singles = [("island 1080p 14h", 5, 14.0, 1920, 1080, -4.5), ("mountains 4K sea 3.5", 6, 14.0, 3840, 2160, 3.5),
           ("mountains 720p 14h", 6, 14.0, 1280, 720, -4.5)]
s = rt.Scene()
for rnd in range(2):                    # two rounds, interleaved: drift shows up as disagreement between them
    for v in values:
        for r, _ in ctxs:
            r.set_option(key, v)
        for _ in range(5):
            flush.fill_(1)
            for r, fr in ctxs:
                r.render(fr, out)
        ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(reps)]
        for k in range(reps):
            flush.fill_(k & 255)
            ev[k][0].record(stream)
            for r, fr in ctxs:
                r.render(fr, out)
            ev[k][1].record(stream)
        torch.cuda.synchronize()
        ms = sum(a.elapsed_time(b) for a, b in ev) / reps
        line = {"key": key, "value": v, "round": rnd, "bench_ms_per_step": ms, "bench_Grays_s": rays / ms / 1e6}
        r0 = ctxs[0][0]
        for name, view, hours, w, h, sea in singles:
            f = s.configure(view=view, hours=hours, imgw=w, imgh=h, sea_level=sea, alias=True)
            r0.set_scene(s.objects, s.lights)
            r0.set_option(_lib.RT_OPT_TIMING, 1)
            for _ in range(3):
                r0.render(f, out)
            tr = fx = 0.0
            for _ in range(reps):
                r0.render(f, out)
                t = r0.timings()
                tr += t["trace_ms"]
                fx += t["fxaa_ms"]
            r0.set_option(_lib.RT_OPT_TIMING, 0)
            line[name] = {"trace_us": 1e3 * tr / reps, "fxaa_us": 1e3 * fx / reps}
        # config 5: 32 frames of the day/night path in order (lights move: refit + light bins rebuilt per frame)
        path = configs.config5(tuple(range(0, 256, 8)))
        for rep in range(2):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(stream)
            for _, objs, lights, f in path:
                r0.set_scene(objs, lights)
                r0.render(f, out)
            e1.record(stream)
            torch.cuda.synchronize()
        line["config5 path us/frame (set_scene + render)"] = 1e3 * e0.elapsed_time(e1) / len(path)
        objs, lights, fr = states[0][1:]
        r0.set_scene(objs, lights)
        print(json.dumps(line), flush=True)
