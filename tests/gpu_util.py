"""Shared helpers of the -m gpu tests: every call below goes through the C-ABI of libraytracing_b200.so."""
import ctypes as C

import numpy as np

import raytracing_cuda_b200 as rt
from raytracing_cuda_b200 import _lib

_state = {}


def sky_small():
    if "sky" not in _state:
        _state["sky"] = rt.synthetic_sky(512, 256)
    return _state["sky"]


def renderer(sky=None):
    """One long-lived context per test process (sky upload is the slow part)."""
    if "r" not in _state:
        r = rt.Renderer(0)
        r.set_sky(sky_small() if sky is None else sky)
        _state["r"] = r
    r = _state["r"]
    for k in (_lib.RT_OPT_TRAVERSAL, _lib.RT_OPT_COUNTERS, _lib.RT_OPT_FXAA_IMPL, _lib.RT_OPT_TIMING,
              _lib.RT_OPT_SKY_ALL4, _lib.RT_OPT_TRACE_IMPL, _lib.RT_OPT_BUILDER, _lib.RT_OPT_FORCE_RADIX,
              _lib.RT_OPT_ORDER, _lib.RT_OPT_SHADOW_BINS, _lib.RT_OPT_MIRROR_BINS, _lib.RT_OPT_LIGHT_BINS_REUSE,
              _lib.RT_OPT_COMPACTION, _lib.RT_OPT_PACKET, _lib.RT_OPT_BIG_STACK, _lib.RT_OPT_WIDE_BVH):
        r.set_option(k, 0)
    return r


def render(r, scene, frame, want_ids=True, shard=None, objects=None, lights=None):
    r.set_scene(scene.objects if objects is None else objects, scene.lights if lights is None else lights)
    return r.render_to_host(frame, want_ids=want_ids, shard=shard)


def to_dev(arr):
    import torch
    return torch.from_numpy(np.ascontiguousarray(arr).view(np.int32)).cuda()


def to_host_u32(t, shape):
    return t.cpu().numpy().view(np.uint32).reshape(shape)
