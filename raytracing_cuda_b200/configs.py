"""The five BASELINE.json configs as data: which scene states, poses and frame sizes each one is.

Shared by the parity tests, bench.py and tools/run_configs.py so that all three render the same frames.
Every state comes out of the host scene layer (scene.py -> csrc/scene_host.cpp), i.e. the restated
reference scene.cpp; nothing here touches a GPU.
"""
import ctypes as C
import os

import numpy as np

from .scene import POSES, TIME_PRESETS, Scene, load_sky_pngs, synthetic_sky, terrain_objects

SKY_W, SKY_H = 8192, 4096                       # size of the reference's maps; the procedural stand-in matches it
SKY_NAMES = ("morning", "day", "evening", "night")     # texture1..4 of reference scene.cpp:626-632

CONFIG3_SEA_LEVELS = (-4.5, -2.5, -0.5, 1.5, 3.5, 5.5, 7.5, 9.5, 11.5)
CONFIG4_CELLS = 708
CONFIG4_POSE = ((324.4, 330.0, -84.0), 141.2, 35.0)   # positive verAngle looks DOWN (reference scene.cpp:114)
CONFIG5_FRAMES = 256
CONFIG5_DT = 0.1875                             # seconds per frame: 0.5 h/s * 0.1875 s * 256 frames = 24 h


def copy_frame(f):
    g = type(f)()
    C.memmove(C.byref(g), C.byref(f), C.sizeof(f))
    return g


def sky_maps(directory=None):
    """(maps uint8[4][h][w][4], label): the four PNGs the reference loads (backgrounds/morning|day|evening|night.png,
    scene.cpp:626-632) from `directory` when they are all there, else the procedural stand-in of the same size."""
    if directory and all(os.path.isfile(os.path.join(directory, n + ".png")) for n in SKY_NAMES):
        return load_sky_pngs(directory), "reference sky maps"
    return synthetic_sky(SKY_W, SKY_H), "synthetic"


def state(scene, f):
    """Snapshot of what launch() hands to launchKernel for the scene's current state."""
    return scene.objects.copy(), scene.lights.copy(), copy_frame(f)


def config1():
    s = Scene()
    f = s.configure(view=6, hours=14.0, imgw=1280, imgh=720, alias=False)
    return state(s, f)


def config2(imgw=1920, imgh=1080):
    """[(time of day, objects, lights, frame)] x 4: Island view, FXAA on."""
    s = Scene()
    out = []
    for tod in ("morning", "afternoon", "evening", "night"):
        f = s.configure(view=5, hours=TIME_PRESETS[tod], imgw=imgw, imgh=imgh, alias=True)
        out.append((tod,) + state(s, f))
    return out


def config3(sea_levels=CONFIG3_SEA_LEVELS, imgw=3840, imgh=2160):
    s = Scene()
    out = []
    for sea in sea_levels:
        f = s.configure(view=6, hours=14.0, imgw=imgw, imgh=imgh, sea_level=sea, alias=True)
        out.append((sea,) + state(s, f))
    return out


def config4(imgw=3840, imgh=2160, cells=CONFIG4_CELLS, alias=True):
    """(objects, lights, frame): the synthetic terrain (2 cells^2 triangles + sea plane + sun/moon)."""
    objs = terrain_objects(cells)
    s = Scene()
    pos, hor, ver = CONFIG4_POSE
    s.set_pose(pos, hor, ver)
    s.set_time(14.0)
    s.set_antialiasing(alias)
    f = s.frame(imgw, imgh)
    return objs, s.lights.copy(), copy_frame(f)


def config5_time(i):
    """Automatic time of reference scene.cpp:695-697 with a fixed deltaTime: t_i = fmodf(6 + 0.5 dt i, 24)."""
    return float(np.fmod(np.float32(6.0) + np.float32(0.5 * CONFIG5_DT * i), np.float32(24.0)))


def config5(frames=None, imgw=1920, imgh=1080):
    """[(i, objects, lights, frame)] of the 256-frame camera path: pose interpolated from the key-5 to the key-6 pose,
    one full day of automatic time, so all four sky maps and all four cross-fades (scene.cpp:788-803) occur."""
    (p5, h5, v5), (p6, h6, v6) = POSES[5], POSES[6]
    s = Scene()
    out = []
    for i in (range(CONFIG5_FRAMES) if frames is None else frames):
        a = i / float(CONFIG5_FRAMES - 1)
        s.set_pose(tuple(p5[k] + a * (p6[k] - p5[k]) for k in range(3)), h5 + a * (h6 - h5), v5 + a * (v6 - v5))
        s.set_time(config5_time(i))
        s.set_antialiasing(True)
        out.append((i,) + state(s, s.frame(imgw, imgh)))
    return out
