#!/usr/bin/env python3
"""The four frames of a bench step (config 2: four contexts, one per time of day) on ONE stream against each context on
its own stream (fork after the L2 flush, join before the step's end event).  One JSON line per mode and round."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

import raytracing_cuda_b200 as rt  # noqa: E402
from raytracing_cuda_b200 import _lib, configs  # noqa: E402

reps = int(sys.argv[1]) if len(sys.argv) > 1 else 60
sky, label = configs.sky_maps(os.path.join(ROOT, "oracle", "_ref", "backgrounds"))
states = configs.config2()
main = torch.cuda.current_stream()
ctxs, outs, streams = [], [], []
for i, (_, objs, lights, fr) in enumerate(states):
    r = rt.Renderer(0)
    if i == 0:
        r.set_sky(sky)
    else:
        r.share_sky(ctxs[0][0])
    r.set_scene(objs, lights)
    ctxs.append((r, fr))
    outs.append(torch.empty(1920 * 1080, dtype=torch.int32, device="cuda"))
    streams.append(torch.cuda.Stream())
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
rays = 0
for (r, fr), o in zip(ctxs, outs):
    r.set_option(_lib.RT_OPT_COUNTERS, 1)
    r.render(fr, o)
    rays += rt.rays_per_frame(r.counters())
    r.set_option(_lib.RT_OPT_COUNTERS, 0)
torch.cuda.synchronize()
ref = [o.clone() for o in outs]


def step(multi):
    if not multi:
        for (r, fr), o in zip(ctxs, outs):
            r.render(fr, o)
        return
    fork = torch.cuda.Event()
    fork.record(main)
    for (r, fr), o, s in zip(ctxs, outs, streams):
        s.wait_event(fork)
        r.render(fr, o)
        done = torch.cuda.Event()
        done.record(s)
        main.wait_event(done)


for rnd in range(2):
    for multi in (False, True):
        for (r, _), s in zip(ctxs, streams):
            r.set_stream(s if multi else main)
        for _ in range(5):
            flush.fill_(1)
            step(multi)
        ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(reps)]
        for k in range(reps):
            flush.fill_(k & 255)
            ev[k][0].record(main)
            step(multi)
            ev[k][1].record(main)
        torch.cuda.synchronize()
        ms = sum(a.elapsed_time(b) for a, b in ev) / reps
        same = all(bool(torch.equal(a, b)) for a, b in zip(outs, ref))
        print(json.dumps({"streams": 4 if multi else 1, "round": rnd, "ms_per_step": ms, "Grays_s": rays / ms / 1e6,
                          "frames_equal": same}), flush=True)
