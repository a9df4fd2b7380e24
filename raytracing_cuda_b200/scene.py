"""Host scene layer: Python face of rtb_scene_* (csrc/scene_host.cpp).

Mirrors the reference's scene.h entry points (reference scene.h:16-20: initScene, animate,
mouseMotion, launch, getTime) with explicit state: `Scene()` == initScene() minus the
texture load, `animate()` == animate() with the key state passed in, `mouse()` ==
mouseMotion(), `launch()` == launch(), `clock()` == getTime().
"""
import ctypes as C

import numpy as np

from . import _lib
from .structs import CAMERA_DTYPE, LIGHT_DTYPE, OBJECT_DTYPE, Frame

# reference scene.cpp:713-728 (keys 1-4) and 736-747 (keys 5/6)
TIME_PRESETS = {"morning": 6.0, "afternoon": 14.0, "evening": 18.0, "night": 1.0}
VIEW_ISLAND, VIEW_MOUNTAINS = 5, 6
POSES = {5: ((-56.0, 2.2, 72.0), 309.0, -7.07), 6: ((324.4, 12.41, -84.0), 141.2, -12.65)}

KEYS = {name: 1 << i for i, name in enumerate(
    ["W", "S", "A", "D", "Q", "E", "SHIFT", "LEFT", "RIGHT", "UP", "DOWN", "1", "2", "3", "4", "5", "6", "P", "O",
     "B", "V"])}


class Scene:
    def __init__(self):
        self._lib = _lib.load()
        self._h = self._lib.rtb_scene_create()
        if not self._h:
            raise MemoryError("rtb_scene_create")

    def close(self):
        if self._h:
            self._lib.rtb_scene_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- state views (no copies: they alias the C++ state) ----
    @property
    def objects(self):
        n = C.c_int()
        p = self._lib.rtb_scene_objects(self._h, C.byref(n))
        buf = (C.c_ubyte * (OBJECT_DTYPE.itemsize * n.value)).from_address(C.addressof(p.contents))
        return np.frombuffer(buf, dtype=OBJECT_DTYPE, count=n.value)

    @property
    def lights(self):
        n = C.c_int()
        p = self._lib.rtb_scene_lights(self._h, C.byref(n))
        buf = (C.c_ubyte * (LIGHT_DTYPE.itemsize * n.value)).from_address(C.addressof(p.contents))
        return np.frombuffer(buf, dtype=LIGHT_DTYPE, count=n.value)

    @property
    def camera(self):
        p = self._lib.rtb_scene_camera(self._h)
        buf = (C.c_ubyte * CAMERA_DTYPE.itemsize).from_address(C.addressof(p.contents))
        return np.frombuffer(buf, dtype=CAMERA_DTYPE, count=1)[0]

    @property
    def sky_vars(self):
        p = self._lib.rtb_scene_sky_vars(self._h)
        return np.ctypeslib.as_array(p, shape=(4,))

    @property
    def ambient(self):
        p = self._lib.rtb_scene_ambient(self._h)
        return np.ctypeslib.as_array(p, shape=(3,))

    @property
    def time(self):
        return self._lib.rtb_scene_time(self._h)

    # ---- controls ----
    def set_time(self, hours):
        self._lib.rtb_scene_set_time(self._h, float(hours))

    def set_view(self, key):
        if self._lib.rtb_scene_set_view(self._h, int(key)) != 0:
            raise ValueError("view key must be 5 (island) or 6 (mountains)")

    def set_pose(self, pos, hor_angle, ver_angle):
        self._lib.rtb_scene_set_pose(self._h, float(pos[0]), float(pos[1]), float(pos[2]), float(hor_angle),
                                     float(ver_angle))

    def update_camera(self, imgw, imgh):
        self._lib.rtb_scene_update_camera(self._h, int(imgw), int(imgh))

    def set_sea_level(self, y):
        self._lib.rtb_scene_set_sea_level(self._h, float(y))

    def set_antialiasing(self, on):
        self._lib.rtb_scene_set_antialiasing(self._h, 1 if on else 0)

    def animate(self, delta_time, keys=()):
        mask = 0
        for k in keys:
            mask |= KEYS[str(k).upper()]
        self._lib.rtb_scene_animate(self._h, float(delta_time), mask)

    def mouse(self, dx, dy):
        self._lib.rtb_scene_mouse(self._h, int(dx), int(dy))

    def clock(self):
        return self._lib.rtb_scene_clock(self._h).decode()

    def frame(self, imgw, imgh):
        """rt_frame for the current state (what launch() hands to launchKernel)."""
        self.update_camera(imgw, imgh)
        f = Frame()
        self._lib.rtb_scene_frame(self._h, int(imgw), int(imgh), C.byref(f))
        return f

    def configure(self, view=5, hours=14.0, imgw=1280, imgh=720, sea_level=None, alias=True):
        """Harness rule of SURVEY 3.2: pose -> aspect -> corners; time -> sky weights -> colours -> lights."""
        self.set_view(view)
        if sea_level is not None:
            self.set_sea_level(sea_level)
        self.set_time(hours)
        self.set_antialiasing(alias)
        return self.frame(imgw, imgh)


def terrain_objects(cells, seed=0x5EED):
    """BASELINE.json config 4: synthetic tessellated terrain as an Object array."""
    lib = _lib.load()
    n = lib.rtb_terrain_build(int(cells), int(seed), None, 0)
    arr = np.zeros(n, dtype=OBJECT_DTYPE)
    got = lib.rtb_terrain_build(int(cells), int(seed), arr.ctypes.data, n)
    if got != n:
        raise RuntimeError("rtb_terrain_build")
    return arr


def _material_ptr(material):
    if material is None:
        return None, None
    m = np.zeros(1, dtype=OBJECT_DTYPE)
    m[0] = material
    return m, m.ctypes.data


def heightmap_objects(heights, x0, x1, z0, z1, material=None):
    """Heightmap [nz][nx] over [x0,x1] x [z0,z1] -> triangle Object array (csrc/ingest.cpp)."""
    lib = _lib.load()
    hm = np.ascontiguousarray(heights, dtype=np.float32)
    nz, nx = hm.shape
    keep, mp = _material_ptr(material)
    n = lib.rtb_heightmap_objects(hm.ctypes.data, nx, nz, x0, x1, z0, z1, mp, None, 0)
    arr = np.zeros(max(n, 0), dtype=OBJECT_DTYPE)
    if n <= 0 or lib.rtb_heightmap_objects(hm.ctypes.data, nx, nz, x0, x1, z0, z1, mp, arr.ctypes.data, n) != n:
        raise ValueError("rtb_heightmap_objects: need at least 2 x 2 samples")
    return arr


def load_obj(text_or_path, material=None):
    """Wavefront OBJ (text, or a path to a file) -> triangle Object array (csrc/ingest.cpp)."""
    import os
    lib = _lib.load()
    if isinstance(text_or_path, str) and "\n" not in text_or_path and os.path.isfile(text_or_path):
        with open(text_or_path, "rb") as fh:
            data = fh.read()
    else:
        data = text_or_path.encode() if isinstance(text_or_path, str) else bytes(text_or_path)
    keep, mp = _material_ptr(material)
    n = lib.rtb_obj_parse(data, len(data), mp, None, 0)
    if n < 0:
        raise ValueError("rtb_obj_parse: malformed face (rc=%d)" % n)
    arr = np.zeros(n, dtype=OBJECT_DTYPE)
    if n and lib.rtb_obj_parse(data, len(data), mp, arr.ctypes.data, n) != n:
        raise ValueError("rtb_obj_parse")
    return arr


def synthetic_sky(w, h):
    """Four procedural RGBA8 sky maps [4][h][w][4] (asset-free stand-in for backgrounds/*.png)."""
    lib = _lib.load()
    out = np.empty((4, h, w, 4), dtype=np.uint8)
    for k in range(4):
        lib.rtb_sky_synthetic(k, int(w), int(h), out[k].ctypes.data)
    return out


def read_png(path):
    """PNG file -> uint8 [h][w][4] RGBA (csrc/image_out.cpp rtb_read_png: what stbi_load(path, ..., 4) returns in the
    reference, scene.cpp:626-632)."""
    lib = _lib.load()
    w, h = C.c_int(), C.c_int()
    p = str(path).encode()
    rc = lib.rtb_read_png(p, C.byref(w), C.byref(h), None, 0)
    if rc != 0:
        raise ValueError("rtb_read_png(%s): rc=%d" % (path, rc))
    out = np.empty((h.value, w.value, 4), dtype=np.uint8)
    rc = lib.rtb_read_png(p, C.byref(w), C.byref(h), out.ctypes.data, out.nbytes)
    if rc != 0:
        raise ValueError("rtb_read_png(%s): rc=%d" % (path, rc))
    return out


def load_sky_pngs(directory):
    """The reference's own sky maps in the order of its texture1..4 (reference scene.cpp:626-632)."""
    import os
    return np.ascontiguousarray(np.stack([read_png(os.path.join(directory, name + ".png"))
                                          for name in ("morning", "day", "evening", "night")]))
