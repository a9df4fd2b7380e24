"""Multi-GPU partitioning of the hot path.  Pure index math + torch.distributed plumbing.

Two ways to use N GPUs, one process per GPU:

* frame-parallel (camera-path batches, BASELINE config 5): frame i -> rank i % N, no communication.
* screen-tile shards of ONE frame (configs 3/4): tile (i, j) of tile_w x tile_h pixels belongs to rank
  (i + j*stride) % N.  A rank traces its tiles plus a 1-pixel apron (rt_shard_needs in
  csrc/trace_common.cuh), so FXAA needs no halo exchange; the only collective is the optional gather of
  finished tiles to rank 0 (`gather_frame`): each rank packs the tiles it owns into a dense buffer
  (rt_pack_tiles), one `dist.gather` moves them over NVLink (NCCL) and rank 0 scatters them into the frame
  (rt_unpack_tiles).  Buffers are equally sized on every rank: slots = ceil(tiles_x / N) * tiles_y.

The numpy pack/unpack below restate the CUDA kernels for the CPU (gloo) tests of this logic.
"""
import ctypes as C

import numpy as np

from . import _lib
from .renderer import _ptr, default_stride, make_shard  # noqa: F401


def tile_grid(imgw, imgh, tile_w, tile_h):
    return (imgw + tile_w - 1) // tile_w, (imgh + tile_h - 1) // tile_h


def owner_map(imgw, imgh, tile_w, tile_h, n_shards, stride):
    """[imgh][imgw] int array: which shard owns each pixel (same formula as rt_tile_owner in csrc/rt_device.h)."""
    yy, xx = np.mgrid[0:imgh, 0:imgw]
    return ((xx // tile_w) + (yy // tile_h) * stride) % n_shards


def needed_map(imgw, imgh, tile_w, tile_h, n_shards, stride, shard_id):
    """Pixels a shard has to TRACE: its own tiles dilated by one pixel (the FXAA apron)."""
    own = owner_map(imgw, imgh, tile_w, tile_h, n_shards, stride) == shard_id
    need = own.copy()
    for dy in (-1, 0, 1):
        for dx in (-1, 0, 1):
            shifted = np.zeros_like(own)
            ys = slice(max(dy, 0), imgh + min(dy, 0))
            yd = slice(max(-dy, 0), imgh + min(-dy, 0))
            xs = slice(max(dx, 0), imgw + min(dx, 0))
            xd = slice(max(-dx, 0), imgw + min(-dx, 0))
            shifted[yd, xd] = own[ys, xs]
            need |= shifted
    return need


def tile_slots(imgw, imgh, shard):
    """Slot s of a shard -> (tx, ty) or None for an empty slot.  Mirrors rt_pack_tiles_kernel."""
    tiles_x, tiles_y = tile_grid(imgw, imgh, shard.tile_w, shard.tile_h)
    per_row = (tiles_x + shard.n_shards - 1) // shard.n_shards
    out = []
    for s in range(per_row * tiles_y):
        ty, m = divmod(s, per_row)
        i0 = (shard.shard_id - ty * shard.stride) % shard.n_shards
        tx = i0 + m * shard.n_shards
        out.append((tx, ty) if tx < tiles_x else None)
    return out


def pack_tiles_host(frame, shard):
    imgh, imgw = frame.shape
    slots = tile_slots(imgw, imgh, shard)
    packed = np.zeros((len(slots), shard.tile_h, shard.tile_w), dtype=frame.dtype)
    for s, t in enumerate(slots):
        if t is None:
            continue
        x0, y0 = t[0] * shard.tile_w, t[1] * shard.tile_h
        blk = frame[y0:y0 + shard.tile_h, x0:x0 + shard.tile_w]
        packed[s, :blk.shape[0], :blk.shape[1]] = blk
    return packed


def unpack_tiles_host(packed, frame, shard):
    imgh, imgw = frame.shape
    for s, t in enumerate(tile_slots(imgw, imgh, shard)):
        if t is None:
            continue
        x0, y0 = t[0] * shard.tile_w, t[1] * shard.tile_h
        h = min(shard.tile_h, imgh - y0)
        w = min(shard.tile_w, imgw - x0)
        frame[y0:y0 + h, x0:x0 + w] = packed[s, :h, :w]
    return frame


def frames_of_rank(n_frames, rank, world):
    """Frame-parallel batches: frame i -> rank i % world."""
    return list(range(rank, n_frames, world))


def gather_frame(dist, rank, world, packed_local, make_like, unpack, dst=0):
    """Gather every rank's packed tiles to `dst` and scatter them into the frame there.

    packed_local : this rank's packed tiles (torch tensor; CUDA for NCCL, CPU for gloo)
    make_like    : () -> empty tensor shaped like packed_local (receive buffers on dst)
    unpack       : (packed_tensor, shard_id) -> None, writes the tiles of shard_id into the frame
    """
    if world == 1:
        unpack(packed_local, 0)
        return
    bufs = [make_like() for _ in range(world)] if rank == dst else None
    dist.gather(packed_local, gather_list=bufs, dst=dst)
    if rank == dst:
        for k in range(world):
            unpack(bufs[k], k)


class PeerFrame:
    """One frame over N GPUs without a gather (rt_peer_frame_*, csrc/peer.cu): the owner rank allocates the
    frame and exports a cudaIpc handle, the other ranks open it, and every rank's sharded render stores its
    tiles straight into the owner's memory over NVLink.  `exchange(handle_or_None) -> handle` moves the 64
    handle bytes from the owner to everybody (torch.distributed.broadcast_object_list, a pipe, ...).

        pf = PeerFrame(renderer, rank, world, w * h, exchange)
        for seq in 1, 2, ...:
            pf.render(frame, shard, seq)        # acquire buffer, trace + FXAA into it, signal
            if rank == pf.owner_rank:
                pf.wait(seq); ...use pf.tensor(seq)...; pf.release(seq)
    """

    def __init__(self, renderer, rank, world, n_pixels, exchange, owner_rank=0, n_buffers=2):
        self.r, self.rank, self.world, self.owner_rank = renderer, rank, world, owner_rank
        self.n_pixels, self.n_buffers = int(n_pixels), int(n_buffers)
        self._lib = renderer._lib
        self._pf = C.c_void_p()
        buf = (C.c_ubyte * _lib.RT_PEER_HANDLE_BYTES)()
        if rank == owner_rank:
            renderer._check(self._lib.rt_peer_frame_create(renderer._h, self.n_pixels, self.n_buffers,
                                                           C.byref(self._pf), buf), "rt_peer_frame_create")
            exchange(bytes(buf))
        else:
            handle = exchange(None)
            buf = (C.c_ubyte * _lib.RT_PEER_HANDLE_BYTES).from_buffer_copy(handle)
            renderer._check(self._lib.rt_peer_frame_open(renderer._h, buf, self.n_pixels, self.n_buffers,
                                                         C.byref(self._pf)), "rt_peer_frame_open")

    def buffer_ptr(self, seq):
        return self._lib.rt_peer_frame_buffer(self._pf, seq)

    def render(self, frame, shard, seq, hit_ids=None, fused=True):
        """This rank's tiles of frame `seq` into the owner's buffer.  fused (default): rt_render_peer -- the frame's
        last kernel waits for the buffer and signals completion itself, three launches in all; else the hand-shake
        as two one-thread kernels around a plain rt_render."""
        r = self.r
        if fused:
            r._check(self._lib.rt_render_peer(r._h, C.byref(frame), C.byref(shard), self._pf, self.rank, seq,
                                              _ptr(hit_ids)), "rt_render_peer")
            return
        r._check(self._lib.rt_peer_frame_acquire(r._h, self._pf, seq), "rt_peer_frame_acquire")
        r.render(frame, self.buffer_ptr(seq), hit_ids=hit_ids, shard=shard)
        r._check(self._lib.rt_peer_frame_signal(r._h, self._pf, self.rank, seq), "rt_peer_frame_signal")

    def wait(self, seq, check=True):
        """Owner: every rank has finished frame `seq` (asynchronous on the context's stream).  check=True also
        synchronises and raises if any wait of the hand-shake gave up -- the frame is invalid then; a caller that
        pipelines frames passes check=False and calls check() before it uses a frame."""
        self.r._check(self._lib.rt_peer_frame_wait(self.r._h, self._pf, self.world, seq), "rt_peer_frame_wait")
        if check:
            self.check()

    def check(self):
        e = self.error()
        if e:
            who = ("the owner gave up waiting for rank %d" % (e - 1)) if e < 0x100 else \
                ("rank %d gave up waiting for the owner's release" % (e - 0x100))
            raise _lib.RtError("peer frame invalid: %s (error word 0x%x); reset_error() re-arms it" % (who, e))

    def release(self, seq):
        self.r._check(self._lib.rt_peer_frame_release(self.r._h, self._pf, seq), "rt_peer_frame_release")

    def set_timeout(self, ms):
        """Give-up time of this process's waits (default 10 s): raise it when ranks start far apart."""
        self.r._check(self._lib.rt_peer_frame_set_timeout(self._pf, int(ms)), "rt_peer_frame_set_timeout")

    def error(self):
        e = C.c_uint(0)
        self.r._check(self._lib.rt_peer_frame_error(self.r._h, self._pf, C.byref(e)), "rt_peer_frame_error")
        return e.value

    def reset_error(self):
        self.r._check(self._lib.rt_peer_frame_reset_error(self.r._h, self._pf), "rt_peer_frame_reset_error")

    def tensor(self, seq):
        """The frame buffer of `seq` as a torch int32 tensor (valid in this process; on the owner it is local)."""
        import torch

        class _Wrap:
            pass
        w = _Wrap()
        w.__cuda_array_interface__ = {"shape": (self.n_pixels,), "typestr": "<i4", "data": (self.buffer_ptr(seq), False),
                                      "version": 2}
        return torch.as_tensor(w, device="cuda")

    def close(self):
        if self._pf:
            self.r._check(self._lib.rt_peer_frame_close(self.r._h, self._pf), "rt_peer_frame_close")
            self._pf = None
