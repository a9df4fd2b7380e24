"""ctypes / numpy mirrors of include/rtb_structs.h and the rt_* PODs of include/rtb_api.h.

Layouts are the reference's (reference structs.h:8-51): Camera 108 B, Object 68 B,
Light 28 B, Ray 24 B.  A ctypes caller has to declare them exactly because
launchKernel takes Camera and float3 BY VALUE (reference kernel.cu:406-411).
"""
import ctypes as C

import numpy as np


class Float3(C.Structure):
    _fields_ = [("x", C.c_float), ("y", C.c_float), ("z", C.c_float)]

    def tuple(self):
        return (self.x, self.y, self.z)


class Camera(C.Structure):
    _fields_ = [("pos", Float3), ("horAngle", C.c_float), ("verAngle", C.c_float), ("fov", C.c_float),
                ("dir", Float3), ("LD", Float3), ("LU", Float3), ("RD", Float3), ("RU", Float3),
                ("left", Float3), ("right", Float3)]


class Object(C.Structure):
    _fields_ = [("type", C.c_int), ("shine", C.c_float), ("specular", C.c_float), ("mirror", C.c_float),
                ("color", Float3), ("pos", Float3), ("size", Float3), ("third", Float3),
                ("light", C.c_ubyte), ("_pad", C.c_ubyte * 3)]


class Light(C.Structure):
    _fields_ = [("pos", Float3), ("color", Float3), ("intensity", C.c_float)]


class Ray(C.Structure):
    _fields_ = [("origin", Float3), ("dir", Float3)]


class Frame(C.Structure):
    """rt_frame"""
    _fields_ = [("cam", Camera), ("ambient", C.c_float * 3), ("skyVars", C.c_float * 4), ("dayTime", C.c_float),
                ("alias", C.c_int), ("imgw", C.c_int), ("imgh", C.c_int)]


class Shard(C.Structure):
    """rt_shard"""
    _fields_ = [("tile_w", C.c_int), ("tile_h", C.c_int), ("n_shards", C.c_int), ("shard_id", C.c_int),
                ("stride", C.c_int)]


assert C.sizeof(Camera) == 108 and C.sizeof(Object) == 68 and C.sizeof(Light) == 28 and C.sizeof(Ray) == 24

SPHERE, PLANE, TRIANGLE = 0, 1, 2
OBJECTS_NUMBER, LIGHTS_NUMBER, MAX_DEPTH = 133, 2, 4

_f3 = np.dtype([("x", "<f4"), ("y", "<f4"), ("z", "<f4")])
OBJECT_DTYPE = np.dtype([("type", "<i4"), ("shine", "<f4"), ("specular", "<f4"), ("mirror", "<f4"),
                         ("color", "<f4", 3), ("pos", "<f4", 3), ("size", "<f4", 3), ("third", "<f4", 3),
                         ("light", "u1"), ("_pad", "u1", 3)])
LIGHT_DTYPE = np.dtype([("pos", "<f4", 3), ("color", "<f4", 3), ("intensity", "<f4")])
CAMERA_DTYPE = np.dtype([("pos", "<f4", 3), ("horAngle", "<f4"), ("verAngle", "<f4"), ("fov", "<f4"),
                         ("dir", "<f4", 3), ("LD", "<f4", 3), ("LU", "<f4", 3), ("RD", "<f4", 3), ("RU", "<f4", 3),
                         ("left", "<f4", 3), ("right", "<f4", 3)])
assert OBJECT_DTYPE.itemsize == 68 and LIGHT_DTYPE.itemsize == 28 and CAMERA_DTYPE.itemsize == 108

# fields that carry meaning (padding bytes after `light` are indeterminate in the reference)
OBJECT_FIELDS = ("type", "shine", "specular", "mirror", "color", "pos", "size", "third", "light")


def objects_equal(a, b):
    """Field-wise bit equality of two OBJECT_DTYPE arrays.  Ignores what the reference leaves
    uninitialised: struct padding, and `third` of spheres and planes (createSphere/createGround never
    write it, reference scene.cpp:177-187, 326-336, and objects come from malloc, scene.cpp:445)."""
    if a.shape != b.shape:
        return False
    for f in OBJECT_FIELDS:
        if f == "third":
            m = a["type"] == TRIANGLE
            if a[f][m].tobytes() != b[f][m].tobytes():
                return False
        elif a[f].tobytes() != b[f].tobytes():
            return False
    return True


def as_array(ptr, n, dtype):
    """numpy view (no copy) of n structs at a ctypes pointer."""
    buf = (C.c_ubyte * (dtype.itemsize * n)).from_address(C.addressof(ptr.contents) if hasattr(ptr, "contents") else ptr)
    return np.frombuffer(buf, dtype=dtype, count=n)
