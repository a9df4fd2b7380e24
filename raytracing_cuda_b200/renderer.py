"""Renderer: Python face of the rt_* handle API (csrc/api.cu), one context per GPU.

Device buffers are torch CUDA tensors (torch is the memory / stream plumbing here, not
the compute path); every kernel that runs is the package's own CUDA.
"""
import ctypes as C

import numpy as np

from . import _lib
from .structs import LIGHT_DTYPE, OBJECT_DTYPE, Frame, Shard


def _ptr(x):
    if x is None:
        return None
    if hasattr(x, "data_ptr"):
        return x.data_ptr()
    if isinstance(x, np.ndarray):
        return x.ctypes.data
    return int(x)


def make_shard(n_shards, shard_id, tile_w=64, tile_h=32, stride=None):
    if stride is None:
        stride = default_stride(n_shards)
    return Shard(tile_w, tile_h, n_shards, shard_id, stride)


def default_stride(n_shards):
    """Row-to-row shift of the tile interleave: coprime with n so columns are not owned by one GPU."""
    if n_shards <= 2:
        return 1
    s = n_shards // 2 + 1
    from math import gcd
    while gcd(s, n_shards) != 1:
        s += 1
    return s


class PinnedFrames:
    """n page-locked host frames of n_pixels uint32 (rt_host_alloc): targets of the asynchronous read-back."""

    def __init__(self, n, n_pixels):
        self._lib = _lib.load()
        self._ptrs = []
        self.frames = []
        for _ in range(n):
            p = C.c_void_p()
            if self._lib.rt_host_alloc(C.byref(p), int(n_pixels) * 4) != _lib.RT_OK:
                self.close()
                raise _lib.RtError("rt_host_alloc(%d bytes) failed" % (int(n_pixels) * 4))
            self._ptrs.append(p)
            buf = (C.c_uint32 * int(n_pixels)).from_address(p.value)
            self.frames.append(np.frombuffer(buf, dtype=np.uint32))

    def __getitem__(self, i):
        return self.frames[i]

    def __len__(self):
        return len(self.frames)

    def close(self):
        self.frames = []
        for p in self._ptrs:
            self._lib.rt_host_free(p)
        self._ptrs = []

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class Renderer:
    def __init__(self, device=0, stream=None, handle=None):
        self._lib = _lib.load()
        if handle is not None:          # a context somebody else owns (rt_legacy_context): never destroyed here
            self._h = C.c_void_p(handle)
            self.device = int(device)
            self._keep = []
            self._borrowed = True
            return
        self._borrowed = False
        h = C.c_void_p()
        rc = self._lib.rt_create(int(device), C.byref(h))
        if rc != _lib.RT_OK:
            raise _lib.RtError("rt_create(device=%d) failed rc=%d: no usable CUDA device; this package has "
                               "no CPU fallback" % (device, rc))
        self._h = h
        self.device = int(device)
        self._keep = []
        if stream is not None:
            self.set_stream(stream)

    def close(self):
        if getattr(self, "_h", None):
            if not self._borrowed:
                self._lib.rt_destroy(self._h)
            self._h = None

    @classmethod
    def legacy(cls):
        """The process-global context behind launchKernel (after its first call), e.g. for its read-back slots."""
        h = _lib.load().rt_legacy_context()
        if not h:
            raise _lib.RtError("launchKernel has not been called yet")
        return cls(handle=h)

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _check(self, rc, what):
        _lib.check(self._h, rc, what)

    def set_option(self, key, value):
        self._check(self._lib.rt_set_option(self._h, int(key), int(value)), "rt_set_option")

    def set_stream(self, stream):
        s = stream.cuda_stream if hasattr(stream, "cuda_stream") else int(stream or 0)
        self._check(self._lib.rt_set_stream(self._h, C.c_void_p(s)), "rt_set_stream")

    def set_sky(self, maps):
        """maps: uint8 array [4][h][w][4] in host memory."""
        maps = np.ascontiguousarray(maps, dtype=np.uint8)
        assert maps.ndim == 4 and maps.shape[0] == 4 and maps.shape[3] == 4
        ptrs = (C.c_void_p * 4)(*[maps[k].ctypes.data for k in range(4)])
        self._check(self._lib.rt_set_sky(self._h, ptrs, maps.shape[2], maps.shape[1]), "rt_set_sky")

    def share_sky(self, owner):
        self._check(self._lib.rt_share_sky(self._h, owner._h), "rt_share_sky")
        self._keep.append(owner)

    def fp32_peak_tflops(self):
        v = C.c_float()
        self._check(self._lib.rt_measure_fp32_peak(self._h, C.byref(v)), "rt_measure_fp32_peak")
        return v.value

    def set_scene(self, objects, lights):
        objects = np.ascontiguousarray(objects, dtype=OBJECT_DTYPE)
        lights = np.ascontiguousarray(lights, dtype=LIGHT_DTYPE)
        self._check(self._lib.rt_set_scene(self._h, objects.ctypes.data, len(objects), lights.ctypes.data,
                                           len(lights)), "rt_set_scene")

    def render(self, frame, out, hit_ids=None, shard=None):
        """Asynchronous render into device buffers (torch tensors or raw pointers)."""
        self._check(self._lib.rt_render(self._h, C.byref(frame), C.byref(shard) if shard is not None else None,
                                        _ptr(out), _ptr(hit_ids)), "rt_render")

    def render_to_host(self, frame, want_ids=False, shard=None, out=None):
        npx = frame.imgw * frame.imgh
        if out is None:
            out = np.empty(npx, dtype=np.uint32)
        ids = np.empty(npx, dtype=np.int32) if want_ids else None
        self._check(self._lib.rt_render_to_host(self._h, C.byref(frame), C.byref(shard) if shard is not None else None,
                                                out.ctypes.data, ids.ctypes.data if want_ids else None),
                    "rt_render_to_host")
        img = out.reshape(frame.imgh, frame.imgw)
        return (img, ids.reshape(frame.imgh, frame.imgw)) if want_ids else img

    def render_to_host_async(self, frame, out_host, slot, shard=None):
        """Render + copy to the (page-locked) host frame `out_host` on the context's copy stream; wait_slot(slot)
        before reading it.  Up to RT_MAX_SLOTS frames in flight."""
        self._check(self._lib.rt_render_to_host_async(self._h, C.byref(frame),
                                                      C.byref(shard) if shard is not None else None, _ptr(out_host),
                                                      int(slot)), "rt_render_to_host_async")

    def readback_async(self, src_dev, n_pixels, out_host, slot):
        self._check(self._lib.rt_readback_async(self._h, _ptr(src_dev), int(n_pixels), _ptr(out_host), int(slot)),
                    "rt_readback_async")

    def wait_slot(self, slot):
        self._check(self._lib.rt_wait_slot(self._h, int(slot)), "rt_wait_slot")

    def render_sequence(self, states, prefix, fmt="png", depth=3, on_frame=None):
        """Frame-sequence dump: states = iterable of (objects, lights, frame); frame i is written to
        <prefix><i:06d>.<fmt> while frames i+1 .. i+depth-1 are being rendered and read back.  Returns the count."""
        depth = max(1, min(int(depth), _lib.RT_MAX_SLOTS))
        pinned = None
        pending = []                                   # (index, slot, w, h)
        n = 0

        def retire():
            i, slot, w, h = pending.pop(0)
            self.wait_slot(slot)
            if on_frame is not None:
                on_frame(i, pinned[slot][:w * h].reshape(h, w))
            rc = self._lib.rtb_write_frame(str(prefix).encode(), i, 1 if fmt == "png" else 0,
                                           pinned[slot].ctypes.data, w, h)
            if rc != _lib.RT_OK:
                raise _lib.RtError("rtb_write_frame(%s, %d) failed rc=%d" % (prefix, i, rc))

        for i, (objs, lights, f) in enumerate(states):
            npx = f.imgw * f.imgh
            if pinned is None or len(pinned[0]) < npx:
                while pending:
                    retire()
                pinned = PinnedFrames(depth, npx)
            if len(pending) == depth:
                retire()
            self.set_scene(objs, lights)
            slot = i % depth
            self.render_to_host_async(f, pinned[slot], slot)
            pending.append((i, slot, f.imgw, f.imgh))
            n += 1
        while pending:
            retire()
        return n

    def base_image_ptr(self):
        p = C.c_void_p()
        self._check(self._lib.rt_get_base_image(self._h, C.byref(p)), "rt_get_base_image")
        return p.value

    def fxaa(self, in_dev, out_dev, imgw, imgh, shard=None):
        self._check(self._lib.rt_fxaa(self._h, _ptr(in_dev), _ptr(out_dev), imgw, imgh,
                                      C.byref(shard) if shard is not None else None), "rt_fxaa")

    def present(self, in_dev, out_w, out_h, factor, out_dev, box=False):
        """The reference's `factor` knob (main.cpp:40-47): (out_w*factor x out_h*factor) frame -> window-size frame,
        point-sampled like its GL_NEAREST blit, or box-averaged (supersampling)."""
        self._check(self._lib.rt_present(self._h, _ptr(in_dev), int(out_w), int(out_h), int(factor), 1 if box else 0,
                                         _ptr(out_dev)), "rt_present")

    def pack_tiles(self, frame_dev, imgw, imgh, shard, packed_dev):
        self._check(self._lib.rt_pack_tiles(self._h, _ptr(frame_dev), imgw, imgh, C.byref(shard), _ptr(packed_dev)),
                    "rt_pack_tiles")

    def unpack_tiles(self, packed_dev, imgw, imgh, shard, frame_dev):
        self._check(self._lib.rt_unpack_tiles(self._h, _ptr(packed_dev), imgw, imgh, C.byref(shard), _ptr(frame_dev)),
                    "rt_unpack_tiles")

    def tile_slots(self, shard, imgw, imgh):
        return self._lib.rt_shard_tile_count(C.byref(shard), imgw, imgh)

    def synchronize(self):
        self._check(self._lib.rt_synchronize(self._h), "rt_synchronize")

    def counters(self):
        buf = (C.c_ulonglong * _lib.RT_CNT_COUNT)()
        self._check(self._lib.rt_get_counters(self._h, buf), "rt_get_counters")
        return {name: int(buf[i]) for i, name in enumerate(_lib.COUNTER_NAMES)}

    def device_errors(self):
        """Error word of the device code (bit 0: a traversal stack ran out); synchronises, reading clears it."""
        e = C.c_uint(0)
        self._check(self._lib.rt_get_device_errors(self._h, C.byref(e)), "rt_get_device_errors")
        return e.value

    def light_bins(self):
        """Diagnostics: the light bins of the last small-scene frame as uint16[4 lights, 6 faces, 64, 64, 8]
        ([..., 0] = count or 0xFFFF, [..., 1:] = leaf offsets), or None."""
        import numpy as np
        a = np.zeros((4, 6, 64, 64, 8), dtype=np.uint16)
        n = self._lib.rt_debug_light_bins(self._h, a.ctypes.data, a.nbytes)
        if n < 0:
            self._check(int(-n), "rt_debug_light_bins")
        return a if n == a.nbytes else None

    def timings(self):
        buf = (C.c_float * 3)()
        self._check(self._lib.rt_get_timings(self._h, buf), "rt_get_timings")
        return {"trace_ms": buf[0], "fxaa_ms": buf[1], "bvh_ms": buf[2]}

    def bvh_info(self):
        a, b, c = C.c_int(), C.c_int(), C.c_int()
        self._check(self._lib.rt_bvh_info(self._h, C.byref(a), C.byref(b), C.byref(c)), "rt_bvh_info")
        return {"prims": a.value, "nodes": b.value, "planes": c.value}

    def bvh_download(self):
        info = self.bvh_info()
        nodes = np.zeros((max(info["nodes"], 0), 16), dtype=np.float32)
        ids = np.zeros(info["prims"], dtype=np.int32)
        keys = np.zeros(info["prims"], dtype=np.uint32)
        self._check(self._lib.rt_bvh_download(self._h, nodes.ctypes.data if len(nodes) else None,
                                              ids.ctypes.data if len(ids) else None,
                                              keys.ctypes.data if len(keys) else None), "rt_bvh_download")
        return nodes, ids, keys

    def sort_pairs(self, keys_dev, vals_dev, n):
        self._check(self._lib.rt_sort_pairs(self._h, _ptr(keys_dev), _ptr(vals_dev), int(n)), "rt_sort_pairs")


def rays_per_frame(counters):
    """The benchmark's unit of work: every trace<d<=4> invocation plus every shadow ray
    the reference constructs (reference kernel.cu:182), whether or not this implementation traced it."""
    return counters["rays_primary"] + counters["rays_mirror"] + counters["rays_shadow"]
