"""Tile / frame partitioner (raytracing_cuda_b200/sharding.py): pure index math on the CPU, and the
gather path with world_size 2 over gloo."""
import ctypes as C
import os
import subprocess
import sys

import numpy as np
import pytest

from raytracing_cuda_b200 import _lib, sharding
from raytracing_cuda_b200.renderer import default_stride, make_shard

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.parametrize("n", [1, 2, 3, 4, 8])
@pytest.mark.parametrize("size", [(1920, 1080), (1280, 720), (3840, 2160), (333, 187)])
def test_every_pixel_owned_exactly_once_and_balanced(n, size):
    w, h = size
    stride = default_stride(n)
    own = sharding.owner_map(w, h, 64, 32, n, stride)
    counts = np.bincount(own.ravel(), minlength=n)
    assert counts.sum() == w * h and (counts > 0).all()
    if w * h >= 1280 * 720:
        assert counts.max() <= 1.15 * counts.min()            # interleave balances the load
    # slots enumerate exactly the owned tiles
    tiles_x, tiles_y = sharding.tile_grid(w, h, 64, 32)
    seen = set()
    for k in range(n):
        sh = make_shard(n, k)
        slots = sharding.tile_slots(w, h, sh)
        assert len(slots) == _lib.load().rt_shard_tile_count(C.byref(sh), w, h)
        for t in slots:
            if t is not None:
                assert t not in seen and own[t[1] * 32, t[0] * 64] == k
                seen.add(t)
    assert len(seen) == tiles_x * tiles_y


def test_stride_is_coprime_so_columns_are_shared():
    for n in range(2, 17):
        s = default_stride(n)
        assert np.gcd(s, n) == 1
        own = sharding.owner_map(64 * n, 32 * n, 64, 32, n, s)[::32, ::64]   # one entry per tile
        for col in range(n):
            assert len(set(own[:, col])) == n                  # every column of tiles visits every shard


def test_apron_is_one_pixel():
    n, w, h = 4, 640, 360
    stride = default_stride(n)
    for k in range(n):
        own = sharding.owner_map(w, h, 64, 32, n, stride) == k
        need = sharding.needed_map(w, h, 64, 32, n, stride, k)
        assert (need | own == need).all()
        extra = need & ~own
        # every extra pixel touches an owned pixel (8-neighbourhood) and nothing further out is needed
        ys, xs = np.nonzero(extra)
        for y, x in zip(ys[::97], xs[::97]):
            assert own[max(y - 1, 0):y + 2, max(x - 1, 0):x + 2].any()
        overdraw = need.sum() / own.sum()
        assert 1.05 < overdraw < 1.12                           # 66x34 / 64x32 = 9.6 %


def test_pack_unpack_roundtrip_host():
    rng = np.random.default_rng(0)
    for (w, h), n in (((640, 360), 3), ((333, 187), 2), ((1920, 1080), 8)):
        frame = rng.integers(0, 1 << 32, size=(h, w), dtype=np.uint32)
        out = np.zeros_like(frame)
        for k in range(n):
            sh = make_shard(n, k)
            out = sharding.unpack_tiles_host(sharding.pack_tiles_host(frame, sh), out, sh)
        assert (out == frame).all()


def test_frame_parallel_assignment():
    for world in (1, 2, 4, 8):
        got = sorted(sum((sharding.frames_of_rank(256, r, world) for r in range(world)), []))
        assert got == list(range(256))


_WORKER = r'''
import os, sys
import numpy as np, torch, torch.distributed as dist
sys.path.insert(0, sys.argv[1])
from raytracing_cuda_b200 import sharding
from raytracing_cuda_b200.renderer import make_shard
dist.init_process_group("gloo")
rank, world = dist.get_rank(), dist.get_world_size()
w, h = 333, 187
truth = (np.arange(w * h, dtype=np.uint32).reshape(h, w) * 2654435761 % (1 << 24)).astype(np.int32)
sh = make_shard(world, rank)
# each rank only knows the pixels it owns
own = sharding.owner_map(w, h, sh.tile_w, sh.tile_h, world, sh.stride) == rank
mine = np.where(own, truth, -1).astype(np.int32)
packed = torch.from_numpy(sharding.pack_tiles_host(mine, sh))
frame = np.zeros((h, w), dtype=np.int32)
def unpack(buf, k):
    sharding.unpack_tiles_host(buf.numpy(), frame, make_shard(world, k))
sharding.gather_frame(dist, rank, world, packed, lambda: torch.empty_like(packed), unpack, dst=0)
if rank == 0:
    assert (frame == truth).all(), "composited frame differs"
    print("GATHER_OK")
dist.barrier()
dist.destroy_process_group()
'''


def test_gather_world_size_2_gloo(tmp_path):
    script = tmp_path / "worker.py"
    script.write_text(_WORKER)
    env = dict(os.environ, MASTER_ADDR="127.0.0.1")
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2",
                        "--master-addr", "127.0.0.1", "--master-port", "29711", str(script), ROOT],
                       capture_output=True, text=True, timeout=240, env=env)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert "GATHER_OK" in r.stdout
