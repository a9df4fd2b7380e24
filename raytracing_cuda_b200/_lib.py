"""Loader of the C-ABI shared library (include/rtb_api.h).

There is no CPU fallback: if libraytracing_b200.so is missing the import fails loudly
and tells the user to build it.  Host-only entry points (rtb_scene_*, rtb_terrain_build,
rtb_sky_synthetic, rtb_write_*) work without a GPU; rt_create reports RT_ERR_CUDA there.
"""
import ctypes as C
import os

from .structs import Camera, Float3, Frame, Light, Object, Shard

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("RTB_LIB_PATH") or os.path.join(_HERE, "libraytracing_b200.so")   # override: kernel experiments

RT_OK, RT_ERR_CUDA, RT_ERR_ARG, RT_ERR_STATE, RT_ERR_UNSUPPORTED = 0, 1, 2, 3, 4
(RT_OPT_TRAVERSAL, RT_OPT_COUNTERS, RT_OPT_FXAA_IMPL, RT_OPT_TIMING, RT_OPT_SKY_ALL4, RT_OPT_TRACE_IMPL) = range(6)
RT_OPT_BUILDER = 6
RT_OPT_FORCE_RADIX = 7
RT_OPT_ORDER = 8
RT_OPT_SHADOW_BINS = 9
RT_OPT_MIRROR_BINS = 10
RT_OPT_LIGHT_BINS_REUSE = 11
RT_OPT_COMPACTION = 12
RT_OPT_PACKET = 13
RT_OPT_BIG_STACK = 14
RT_OPT_WIDE_BVH = 15
RT_PEER_HANDLE_BYTES = 64
RT_MAX_SLOTS = 8
RT_SORT_STATIC_MAX = 1 << 20
RT_CNT_COUNT = 16
COUNTER_NAMES = ("rays_primary", "rays_shadow", "rays_mirror", "node_tests", "tri_tests", "sphere_tests",
                 "plane_tests", "sky_lookups", "shadow_skipped", "tri_exit_u", "tri_exit_v", "tri_full", "tri_accept",
                 "sphere_hits")

# every symbol include/rtb_api.h declares: name -> (restype, argtypes)
_P = C.c_void_p
_u8p = C.POINTER(C.c_ubyte)
SIGNATURES = {
    "launchKernel": (None, [_P, C.c_int, C.c_int, Camera, C.POINTER(Object), C.POINTER(Light), Float3,
                            C.POINTER(C.c_float), _P, _P, _P, _P, C.c_int, C.c_int, C.c_float, C.c_bool]),
    "rtb_legacy_shutdown": (C.c_int, []),
    "rt_create": (C.c_int, [C.c_int, C.POINTER(_P)]),
    "rt_destroy": (C.c_int, [_P]),
    "rt_last_error": (C.c_char_p, [_P]),
    "rt_set_option": (C.c_int, [_P, C.c_int, C.c_int]),
    "rt_set_stream": (C.c_int, [_P, _P]),
    "rt_set_sky": (C.c_int, [_P, C.POINTER(_P), C.c_int, C.c_int]),
    "rt_share_sky": (C.c_int, [_P, _P]),
    "rt_measure_fp32_peak": (C.c_int, [_P, C.POINTER(C.c_float)]),
    "rt_set_scene": (C.c_int, [_P, _P, C.c_int, _P, C.c_int]),
    "rt_render": (C.c_int, [_P, C.POINTER(Frame), C.POINTER(Shard), _P, _P]),
    "rt_render_to_host": (C.c_int, [_P, C.POINTER(Frame), C.POINTER(Shard), _P, _P]),
    "rt_render_to_host_async": (C.c_int, [_P, C.POINTER(Frame), C.POINTER(Shard), _P, C.c_int]),
    "rt_readback_async": (C.c_int, [_P, _P, C.c_size_t, _P, C.c_int]),
    "rt_wait_slot": (C.c_int, [_P, C.c_int]),
    "rt_host_alloc": (C.c_int, [C.POINTER(_P), C.c_size_t]),
    "rt_host_free": (C.c_int, [_P]),
    "rt_legacy_context": (_P, []),
    "rt_get_base_image": (C.c_int, [_P, C.POINTER(_P)]),
    "rt_fxaa": (C.c_int, [_P, _P, _P, C.c_int, C.c_int, C.POINTER(Shard)]),
    "rt_shard_tile_count": (C.c_int, [C.POINTER(Shard), C.c_int, C.c_int]),
    "rt_pack_tiles": (C.c_int, [_P, _P, C.c_int, C.c_int, C.POINTER(Shard), _P]),
    "rt_unpack_tiles": (C.c_int, [_P, _P, C.c_int, C.c_int, C.POINTER(Shard), _P]),
    "rt_peer_frame_create": (C.c_int, [_P, C.c_size_t, C.c_int, C.POINTER(_P), _u8p]),
    "rt_peer_frame_open": (C.c_int, [_P, _u8p, C.c_size_t, C.c_int, C.POINTER(_P)]),
    "rt_peer_frame_buffer": (_P, [_P, C.c_uint]),
    "rt_peer_frame_acquire": (C.c_int, [_P, _P, C.c_uint]),
    "rt_peer_frame_signal": (C.c_int, [_P, _P, C.c_int, C.c_uint]),
    "rt_peer_frame_wait": (C.c_int, [_P, _P, C.c_int, C.c_uint]),
    "rt_peer_frame_release": (C.c_int, [_P, _P, C.c_uint]),
    "rt_render_peer": (C.c_int, [_P, C.POINTER(Frame), C.POINTER(Shard), _P, C.c_int, C.c_uint, _P]),
    "rt_peer_frame_set_timeout": (C.c_int, [_P, C.c_uint]),
    "rt_peer_frame_error": (C.c_int, [_P, _P, C.POINTER(C.c_uint)]),
    "rt_peer_frame_reset_error": (C.c_int, [_P, _P]),
    "rt_peer_frame_close": (C.c_int, [_P, _P]),
    "rt_present": (C.c_int, [_P, _P, C.c_int, C.c_int, C.c_int, C.c_int, _P]),
    "rt_synchronize": (C.c_int, [_P]),
    "rt_get_counters": (C.c_int, [_P, C.POINTER(C.c_ulonglong)]),
    "rt_debug_light_bins": (C.c_longlong, [_P, C.c_void_p, C.c_size_t]),
    "rt_get_device_errors": (C.c_int, [_P, C.POINTER(C.c_uint)]),
    "rt_launch_count": (C.c_ulonglong, []),
    "rt_get_timings": (C.c_int, [_P, C.POINTER(C.c_float)]),
    "rt_bvh_info": (C.c_int, [_P, C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(C.c_int)]),
    "rt_bvh_download": (C.c_int, [_P, _P, _P, _P]),
    "rt_sort_pairs": (C.c_int, [_P, _P, _P, C.c_int]),
    "rt_sort_pairs_static": (C.c_int, [_P, _P, _P, C.c_int]),
    "rtb_scene_create": (_P, []),
    "rtb_scene_destroy": (None, [_P]),
    "rtb_scene_objects": (C.POINTER(Object), [_P, C.POINTER(C.c_int)]),
    "rtb_scene_lights": (C.POINTER(Light), [_P, C.POINTER(C.c_int)]),
    "rtb_scene_camera": (C.POINTER(Camera), [_P]),
    "rtb_scene_sky_vars": (C.POINTER(C.c_float), [_P]),
    "rtb_scene_ambient": (C.POINTER(C.c_float), [_P]),
    "rtb_scene_time": (C.c_float, [_P]),
    "rtb_scene_set_time": (None, [_P, C.c_float]),
    "rtb_scene_set_view": (C.c_int, [_P, C.c_int]),
    "rtb_scene_set_pose": (None, [_P, C.c_float, C.c_float, C.c_float, C.c_float, C.c_float]),
    "rtb_scene_update_camera": (None, [_P, C.c_int, C.c_int]),
    "rtb_scene_set_sea_level": (None, [_P, C.c_float]),
    "rtb_scene_set_antialiasing": (None, [_P, C.c_int]),
    "rtb_scene_animate": (None, [_P, C.c_float, C.c_uint]),
    "rtb_scene_mouse": (None, [_P, C.c_int, C.c_int]),
    "rtb_scene_clock": (C.c_char_p, [_P]),
    "rtb_scene_frame": (None, [_P, C.c_int, C.c_int, C.POINTER(Frame)]),
    "rtb_scene_launch": (None, [_P, _P, C.c_int, C.c_int, C.POINTER(_P), C.c_int, C.c_int]),
    "rtb_terrain_build": (C.c_int, [C.c_int, C.c_uint, _P, C.c_int]),
    "rtb_heightmap_objects": (C.c_int, [_P, C.c_int, C.c_int, C.c_float, C.c_float, C.c_float, C.c_float, _P, _P, C.c_int]),
    "rtb_obj_parse": (C.c_int, [C.c_char_p, C.c_size_t, _P, _P, C.c_int]),
    "rtb_sky_synthetic": (None, [C.c_int, C.c_int, C.c_int, _P]),
    "rtb_write_ppm": (C.c_int, [C.c_char_p, _P, C.c_int, C.c_int]),
    "rtb_write_png": (C.c_int, [C.c_char_p, _P, C.c_int, C.c_int]),
    "rtb_write_frame": (C.c_int, [C.c_char_p, C.c_int, C.c_int, _P, C.c_int, C.c_int]),
    "rtb_read_png": (C.c_int, [C.c_char_p, C.POINTER(C.c_int), C.POINTER(C.c_int), _P, C.c_size_t]),
}

_lib = None


class LibraryMissing(ImportError):
    pass


def load():
    """dlopen the product library and bind every declared symbol.  Raises if it is not built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.isfile(LIB_PATH):
        raise LibraryMissing(
            "%s is not built: run `python -m raytracing_cuda_b200.build` (needs nvcc). "
            "This package has no CPU fallback." % LIB_PATH)
    lib = C.CDLL(LIB_PATH, mode=C.RTLD_LOCAL)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)   # AttributeError if the symbol is missing
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


class RtError(RuntimeError):
    pass


def check(ctx, rc, what):
    if rc != RT_OK:
        msg = load().rt_last_error(ctx) if ctx else b""
        raise RtError("%s failed: rc=%d %s" % (what, rc, (msg or b"").decode(errors="replace")))
