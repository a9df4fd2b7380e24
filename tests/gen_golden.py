#!/usr/bin/env python3
"""Generate tests/golden/*.npz from the REFERENCE's own code (run in the build container).

Needs /root/reference (through oracle/build_ref.py -> oracle/_ref/*.so):
  - scene_states.npz : Object[133] / Light[2] / Camera / skyVars / ambient produced by the reference's
                       scene.cpp for both views x 10 times of day x 2 sea levels, plus a 40-tick
                       animate() run (time advancing, colour lag) -- reference scene.cpp:100-126, 444-488,
                       666-687, 758-816
  - frames.npz       : images, pre-FXAA images and primary hit ids rendered by the reference's kernel.cu
                       text compiled for the host (oracle/_ref/libref_trace_cpu.so) at 320x180 for both
                       views x 4 times of day, FXAA on, with the package's procedural 512x256 sky maps
                       (the reference's PNG assets do not travel to the GPU box)
  - fxaa.npz         : the reference FXAA kernel text on a random-noise and a hard-edge image
The committed files pin the oracle (tests/test_oracle.py) and the host scene layer
(tests/test_scene_host.py) wherever /root/reference is absent.
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
import oracle_util as ou  # noqa: E402

sys.path.insert(0, os.path.join(ou.ROOT, "oracle"))
import build_ref  # noqa: E402

import raytracing_cuda_b200 as rt  # noqa: E402

TIMES = [6.0, 14.0, 18.0, 1.0, 9.3, 17.5, 21.1, 4.7, 0.0, 23.99]
SEAS = [-4.5, 3.5]
FRAME_W, FRAME_H = 320, 180
SKY_W, SKY_H = 512, 256


def state(ref):
    return dict(objects=ref.objects(), lights=ref.lights(), camera=np.array([ref.camera()]), sky=ref.sky_vars(),
                ambient=ref.ambient())


def main():
    build_ref.build_all(gpu_ref=False)
    os.makedirs(ou.GOLDEN, exist_ok=True)
    ref = ou.RefScene()
    out = {}
    case = 0
    meta = []
    for view in (5, 6):
        pos, hor, ver = rt.POSES[view]
        for hours in TIMES:
            for sea in SEAS:
                ref.lib.ref_scene_set_aspect(1920, 1080)
                ref.lib.ref_scene_set_pose(*pos, hor, ver)
                ref.lib.ref_scene_set_sea(sea)
                ref.lib.ref_scene_set_time(hours)
                for k, v in state(ref).items():
                    out["c%03d_%s" % (case, k)] = v
                meta.append((view, hours, sea))
                case += 1
    out["meta"] = np.array(meta, dtype=np.float64)
    # animate(): 40 ticks of 0.75 s from 05:00 with play on (crosses the 4-6 h and 6 h boundaries)
    ref.lib.ref_scene_set_sea(-4.5)
    ref.lib.ref_scene_set_time(5.0)
    ref.lib.ref_scene_set_play(1)
    for tick in range(40):
        ref.lib.ref_scene_animate(0.75)
        for k, v in state(ref).items():
            out["a%03d_%s" % (tick, k)] = v
        out["a%03d_time" % tick] = np.float32(ref.lib.ref_scene_time())
        out["a%03d_clock" % tick] = np.frombuffer(ref.lib.ref_scene_clock(), dtype=np.uint8)
    np.savez_compressed(os.path.join(ou.GOLDEN, "scene_states.npz"), **out)

    sky = rt.synthetic_sky(SKY_W, SKY_H)
    s = rt.Scene()
    frames = {}
    for view in (5, 6):
        for name, hours in rt.TIME_PRESETS.items():
            f = s.configure(view=view, hours=hours, imgw=FRAME_W, imgh=FRAME_H, alias=True)
            r = ou.ref_cpu_render(s.objects, s.lights, f, sky)
            key = "v%d_%s" % (view, name)
            frames[key + "_out"] = r["out"]
            frames[key + "_base"] = r["base"]
            frames[key + "_ids"] = r["ids"].astype(np.int16)
    np.savez_compressed(os.path.join(ou.GOLDEN, "frames.npz"), **frames)

    # FXAA alone through the reference kernel text: feed an image as "base" by rendering nothing --
    # the wrapper renders first, so instead run the oracle-independent path: craft frames whose base we
    # know by reusing ref_cpu_render's base output and checking out==fxaa(base) is done in tests; here we
    # store two synthetic inputs and the reference FXAA of them via a dedicated entry point.
    rng = np.random.default_rng(7)
    noise = rng.integers(0, 1 << 24, size=(70, 100), dtype=np.uint32)
    edges = np.zeros((70, 100), dtype=np.uint32)
    edges[:, 50:] = 0x00FFFFFF
    edges[20:40, 10:30] = 0x000000FF
    edges[np.add.outer(np.arange(70), np.arange(100)) % 17 == 0] = 0x0000FF00
    fx = {"noise_in": noise, "edges_in": edges}
    lib = ou.ref_lib("trace_cpu")
    if hasattr(lib, "ref_cpu_fxaa"):
        import ctypes as C
        for name in ("noise", "edges"):
            src = fx[name + "_in"]
            dst = np.zeros_like(src)
            lib.ref_cpu_fxaa(C.c_void_p(src.ctypes.data), C.c_void_p(dst.ctypes.data), C.c_int(src.shape[1]),
                             C.c_int(src.shape[0]))
            fx[name + "_out"] = dst
    np.savez_compressed(os.path.join(ou.GOLDEN, "fxaa.npz"), **fx)
    for fn in sorted(os.listdir(ou.GOLDEN)):
        print(fn, os.path.getsize(os.path.join(ou.GOLDEN, fn)))


if __name__ == "__main__":
    main()
