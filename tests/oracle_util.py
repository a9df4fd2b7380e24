"""ctypes access to the oracle (oracle/liboracle.so) and, when present, the reference builds
under oracle/_ref.  TEST INFRASTRUCTURE: only tests/, smoke() and bench.py's CPU legs import this."""
import ctypes as C
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))

from raytracing_cuda_b200.structs import (CAMERA_DTYPE, LIGHT_DTYPE, OBJECT_DTYPE, Camera, Float3,  # noqa: E402
                                          Light, Object)

ORACLE_SO = os.path.join(ROOT, "oracle", "liboracle.so")
REF_DIR = os.path.join(ROOT, "oracle", "_ref")
GOLDEN = os.path.join(ROOT, "tests", "golden")
# the reference's own sky maps (scene.cpp:626-632), copied next to the reference builds by oracle/build_ref.py
REF_BACKGROUNDS = os.path.join(REF_DIR, "backgrounds")

_cache = {}


def oracle_lib():
    if "oracle" not in _cache:
        if not os.path.isfile(ORACLE_SO):
            import build_ref
            build_ref.build_oracle()
        lib = C.CDLL(ORACLE_SO)
        lib.oracle_render.restype = C.c_int
        lib.oracle_fxaa.restype = C.c_int
        lib.oracle_check_hit.restype = C.c_int
        _cache["oracle"] = lib
    return _cache["oracle"]


def ref_lib(name):
    """name in {scene, trace_cpu, kernel_sm100}; None when that reference build is absent."""
    key = "ref_" + name
    if key not in _cache:
        path = os.path.join(REF_DIR, "libref_%s.so" % name)
        _cache[key] = C.CDLL(path, mode=C.RTLD_LOCAL) if os.path.isfile(path) else None
    return _cache[key]


def _frame_args(frame):
    cam = Camera()
    C.memmove(C.byref(cam), C.byref(frame.cam), C.sizeof(Camera))
    amb = (C.c_float * 3)(*frame.ambient)
    sky = (C.c_float * 4)(*frame.skyVars)
    return cam, amb, sky


def oracle_render(objects, lights, frame, sky_maps, x_step=1, y_step=1, want_out=True):
    """Returns dict(out, base, ids, counters) from the repo's CPU restatement."""
    lib = oracle_lib()
    objects = np.ascontiguousarray(objects, dtype=OBJECT_DTYPE)
    lights = np.ascontiguousarray(lights, dtype=LIGHT_DTYPE)
    w, h = frame.imgw, frame.imgh
    cam, amb, sky = _frame_args(frame)
    tex = (C.c_void_p * 4)(*[sky_maps[k].ctypes.data for k in range(4)])
    out = np.zeros((h, w), dtype=np.uint32)
    base = np.zeros((h, w), dtype=np.uint32)
    ids = np.full((h, w), -2, dtype=np.int32)
    cnt = (C.c_ulonglong * 5)()
    rc = lib.oracle_render(C.c_void_p(objects.ctypes.data), C.c_int(len(objects)), C.c_void_p(lights.ctypes.data),
                           C.c_int(len(lights)), C.byref(cam), amb, sky, tex, C.c_int(sky_maps.shape[2]),
                           C.c_int(sky_maps.shape[1]), C.c_float(frame.dayTime), C.c_int(frame.alias), C.c_int(w),
                           C.c_int(h), C.c_int(x_step), C.c_int(y_step),
                           C.c_void_p(out.ctypes.data) if want_out else None, C.c_void_p(base.ctypes.data),
                           C.c_void_p(ids.ctypes.data), cnt)
    assert rc == 0
    names = ("rays_primary", "rays_shadow", "rays_mirror", "prim_tests", "sky_lookups")
    return {"out": out, "base": base, "ids": ids, "counters": {n: int(cnt[i]) for i, n in enumerate(names)}}


def oracle_fxaa(img):
    lib = oracle_lib()
    img = np.ascontiguousarray(img, dtype=np.uint32)
    out = np.empty_like(img)
    lib.oracle_fxaa(C.c_void_p(img.ctypes.data), C.c_void_p(out.ctypes.data), C.c_int(img.shape[1]),
                    C.c_int(img.shape[0]))
    return out


def _check_hit(fn, obj, origin, direction):
    o = np.ascontiguousarray(obj, dtype=OBJECT_DTYPE).reshape(1)
    org = (C.c_float * 3)(*origin)
    d = (C.c_float * 3)(*direction)
    pos, nrm, t = (C.c_float * 3)(), (C.c_float * 3)(), C.c_float()
    hit = fn(C.c_void_p(o.ctypes.data), org, d, pos, nrm, C.byref(t))
    return (bool(hit), t.value, tuple(pos), tuple(nrm))


def oracle_check_hit(obj, origin, direction):
    return _check_hit(oracle_lib().oracle_check_hit, obj, origin, direction)


def near_tie_report(objects, frame, ids_a, ids_b, radii=(1.0 / 256, 1.0 / 64, 1.0 / 16, 1.0 / 4)):
    """For every pixel where two primary-hit-id buffers differ (a = the implementation under test, b = the CPU oracle):
    (x, y, id_a, id_b, t_a, t_b, flip) with the distances the oracle's checkHit gives both objects along the oracle's
    primary ray (nan: not hit), and `flip` = the smallest probe radius in pixels at which the oracle ITSELF, choosing
    between just these two objects, picks id_a for one of the eight rays around the pixel centre (inf: never).
    A small flip radius is what a near tie is: the pixel centre sits on the boundary between the two answers."""
    lib = oracle_lib()
    lib.oracle_primary_ray.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_float, C.c_float, C.c_void_p, C.c_void_p]
    cam = Camera()
    C.memmove(C.byref(cam), C.byref(frame.cam), C.sizeof(Camera))
    org, d = (C.c_float * 3)(), (C.c_float * 3)()

    def winner(px, py, ia, ib):
        lib.oracle_primary_ray(C.byref(cam), frame.imgw, frame.imgh, px, py, org, d)
        best, best_t, ts = -1, float("inf"), []
        for i in sorted({ia, ib}):                       # increasing index, strict '<': the reference's tie rule
            hit, t = (False, 0.0) if i < 0 else oracle_check_hit(objects[i], tuple(org), tuple(d))[:2]
            ts.append((i, t if hit else float("nan")))
            if hit and t < best_t:
                best, best_t = i, t
        return best, dict(ts)

    rows = []
    for y, x in np.argwhere(ids_a != ids_b):
        ia, ib = int(ids_a[y, x]), int(ids_b[y, x])
        _, ts = winner(float(x), float(y), ia, ib)
        flip = float("inf")
        for rad in radii:
            probes = [(dx * rad, dy * rad) for dx in (-1, 0, 1) for dy in (-1, 0, 1) if dx or dy]
            if any(winner(x + dx, y + dy, ia, ib)[0] == ia for dx, dy in probes):
                flip = rad
                break
        rows.append((int(x), int(y), ia, ib, ts.get(ia, float("nan")), ts.get(ib, float("nan")), flip))
    return rows


def ref_check_hit(obj, origin, direction):
    lib = ref_lib("trace_cpu")
    lib.ref_cpu_check_hit.restype = C.c_int
    return _check_hit(lib.ref_cpu_check_hit, obj, origin, direction)


def ref_cpu_render(objects, lights, frame, sky_maps, y_step=1):
    """The reference's own kernel text on the CPU (133 objects exactly)."""
    lib = ref_lib("trace_cpu")
    objects = np.ascontiguousarray(objects, dtype=OBJECT_DTYPE)
    lights = np.ascontiguousarray(lights, dtype=LIGHT_DTYPE)
    assert len(objects) == 133 and len(lights) == 2
    w, h = frame.imgw, frame.imgh
    cam, amb, sky = _frame_args(frame)
    tex = (C.c_void_p * 4)(*[sky_maps[k].ctypes.data for k in range(4)])
    out = np.zeros((h, w), dtype=np.uint32)
    base = np.zeros((h, w), dtype=np.uint32)
    ids = np.full((h, w), -2, dtype=np.int32)
    lib.ref_cpu_render.restype = C.c_int
    rc = lib.ref_cpu_render(C.c_void_p(objects.ctypes.data), C.c_void_p(lights.ctypes.data), C.byref(cam), amb, sky,
                            tex, C.c_int(sky_maps.shape[2]), C.c_int(sky_maps.shape[1]), C.c_float(frame.dayTime),
                            C.c_int(frame.alias), C.c_int(w), C.c_int(h), C.c_int(y_step),
                            C.c_void_p(out.ctypes.data) if y_step == 1 else None,
                            C.c_void_p(base.ctypes.data), C.c_void_p(ids.ctypes.data))
    assert rc == 0
    return {"out": out, "base": base, "ids": ids}


class RefScene:
    """The reference's own scene.cpp (globals!) behind oracle/_ref/libref_scene.so."""

    def __init__(self):
        lib = ref_lib("scene")
        assert lib is not None
        self.lib = lib
        lib.ref_scene_camera.restype = C.POINTER(Camera)
        lib.ref_scene_objects.restype = C.POINTER(Object)
        lib.ref_scene_lights.restype = C.POINTER(Light)
        lib.ref_scene_sky_vars.restype = C.POINTER(C.c_float)
        lib.ref_scene_ambient.restype = C.POINTER(C.c_float)
        lib.ref_scene_time.restype = C.c_float
        lib.ref_scene_clock.restype = C.c_char_p
        lib.ref_scene_set_time.argtypes = [C.c_float]
        lib.ref_scene_animate.argtypes = [C.c_float]
        lib.ref_scene_set_pose.argtypes = [C.c_float] * 5
        lib.ref_scene_set_sea.argtypes = [C.c_float]
        devnull = os.open(os.devnull, os.O_WRONLY)   # initObjects printf("OBJECTS: ...")
        saved = os.dup(1)
        os.dup2(devnull, 1)
        try:
            lib.ref_scene_init()
        finally:
            os.dup2(saved, 1)
            os.close(devnull)
            os.close(saved)

    def objects(self):
        p = self.lib.ref_scene_objects()
        buf = (C.c_ubyte * (68 * 133)).from_address(C.addressof(p.contents))
        return np.frombuffer(buf, dtype=OBJECT_DTYPE, count=133).copy()

    def lights(self):
        p = self.lib.ref_scene_lights()
        buf = (C.c_ubyte * (28 * 2)).from_address(C.addressof(p.contents))
        return np.frombuffer(buf, dtype=LIGHT_DTYPE, count=2).copy()

    def camera(self):
        p = self.lib.ref_scene_camera()
        buf = (C.c_ubyte * 108).from_address(C.addressof(p.contents))
        return np.frombuffer(buf, dtype=CAMERA_DTYPE, count=1).copy()[0]

    def sky_vars(self):
        return np.ctypeslib.as_array(self.lib.ref_scene_sky_vars(), shape=(4,)).copy()

    def ambient(self):
        return np.ctypeslib.as_array(self.lib.ref_scene_ambient(), shape=(3,)).copy()


# ---- image comparison: the gates of BASELINE.json's north_star ----

def unpack_rgb(img):
    img = np.asarray(img, dtype=np.uint32)
    return np.stack([img & 255, (img >> 8) & 255, (img >> 16) & 255], axis=-1).astype(np.int32)


def compare_images(a, b):
    """Returns dict(frac_within_1lsb (per pixel, all channels), psnr_db, max_abs)."""
    ra, rb = unpack_rgb(a), unpack_rgb(b)
    diff = np.abs(ra - rb)
    within = (diff.max(axis=-1) <= 1).mean()
    mse = (diff.astype(np.float64) ** 2).mean()
    psnr = float("inf") if mse == 0 else 10.0 * np.log10(255.0 ** 2 / mse)
    return {"frac_within_1lsb": float(within), "psnr_db": float(psnr), "max_abs": int(diff.max()),
            "frac_exact": float((diff.max(axis=-1) == 0).mean())}
