"""GPU tests of the FXAA kernels, the screen-tile shards and the LBVH builder (all through the C-ABI)."""
import os

import numpy as np
import pytest

import gpu_util as gu
import oracle_util as ou
import raytracing_cuda_b200 as rt
from raytracing_cuda_b200 import _lib, sharding

pytestmark = pytest.mark.gpu


# ---------------------------------------------------------------- FXAA (reference kernel.cu:262-403)

def _fxaa(r, img, impl, shard=None):
    import torch
    h, w = img.shape
    src = gu.to_dev(img)
    dst = torch.zeros_like(src)
    r.set_option(_lib.RT_OPT_FXAA_IMPL, impl)
    r.fxaa(src, dst, w, h, shard)
    r.synchronize()
    return gu.to_host_u32(dst, (h, w))


def _ref_fxaa_gpu(img):
    """The reference's antialiasing kernel rebuilt for sm_100, on a device copy of img (with slack around
    it: the reference's halo loads run out of bounds at the image edges, reference kernel.cu:318-319)."""
    import ctypes as C

    import torch
    lib = ou.ref_lib("kernel_sm100")
    if lib is None:
        return None
    lib.ref_fxaa.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int]
    h, w = img.shape
    pad = w + 4
    big = torch.zeros(w * h + 2 * pad, dtype=torch.int32, device="cuda")
    big[pad:pad + w * h] = gu.to_dev(img).reshape(-1)
    dst = torch.zeros(w * h, dtype=torch.int32, device="cuda")
    lib.ref_fxaa(big[pad:].data_ptr(), dst.data_ptr(), w, h)
    torch.cuda.synchronize()
    return gu.to_host_u32(dst, (h, w)).copy()


def _fxaa_vs_cpu(got, want, noise):
    """CPU (IEEE divides, no FMA) vs GPU (--use_fast_math shapes of the reference): 1 LSB steps where a
    blended channel lands next to an integer, nothing larger except where a near-tie flips the edge choice."""
    m = ou.compare_images(got, want)
    # white noise is the worst case: every pixel is an "edge" and near-ties flip the neighbour choice
    # (measured on B200: 95.5-97 % exact); bit-exactness is asserted against the reference KERNEL instead
    assert m["frac_exact"] >= (0.93 if noise else 0.995), m
    assert m["frac_within_1lsb"] >= (0.94 if noise else 0.999), m


@pytest.mark.parametrize("impl", [0, 1])
def test_fxaa_vs_reference_goldens_and_reference_kernel(impl):
    FX = np.load(os.path.join(ou.GOLDEN, "fxaa.npz"))
    r = gu.renderer()
    for name in ("noise", "edges"):
        got = _fxaa(r, FX[name + "_in"], impl)
        _fxaa_vs_cpu(got, FX[name + "_out"], name == "noise")
        ref = _ref_fxaa_gpu(FX[name + "_in"])
        if ref is not None:      # identical input, same GPU arithmetic: bit for bit
            assert (got == ref).all(), (name, impl, float((got == ref).mean()))


@pytest.mark.parametrize("shape", [(2, 4), (3, 8), (33, 64), (34, 68), (97, 131), (108, 192), (720, 1280), (1080, 1920),
                                   (1088, 2240), (1440, 2560), (2160, 3840)])
def test_fxaa_tma_equals_smem_equals_oracle(shape):
    """Ragged sizes: tiny images, sizes around the 64x32 tile, widths not divisible by 4 (TMA falls back).  Up to
    1,184 tiles (148 SMs x 8) the TMA kernel runs one tile per CTA (1080x1920: 1,020), beyond that CTAs loop over their
    tiles with two stages (1088x2240: 1,190 tiles, just over; 1440x2560; 2160x3840)."""
    rng = np.random.default_rng(shape[0] * 7919 + shape[1])
    h, w = shape
    img = rng.integers(0, 1 << 24, size=(h, w), dtype=np.uint32)
    img[h // 3: h // 2, w // 4: w // 2] = 0x00203040          # a flat patch: pass-through region
    r = gu.renderer()
    a = _fxaa(r, img, 0)
    b = _fxaa(r, img, 1)
    assert (a == b).all()
    _fxaa_vs_cpu(a, ou.oracle_fxaa(img), True)
    ref = _ref_fxaa_gpu(img)
    if ref is not None:
        assert (a == ref).all()
    # borders pass through
    assert (a[0] == img[0]).all() and (a[-1] == img[-1]).all() and (a[:, 0] == img[:, 0]).all() and (a[:, -1] == img[:, -1]).all()


# ---------------------------------------------------------------- shards (multi-GPU partition on one GPU)

@pytest.mark.parametrize("n", [2, 3, 4, 8])
@pytest.mark.parametrize("tile", [(64, 32), (48, 24), (128, 64), (192, 32), (256, 128)])
def test_shard_count_invariance(n, tile):
    """N logical shards on one GPU: every pixel owned exactly once and the composite equals the
    unsharded frame bit for bit (1-pixel apron -> no halo exchange)."""
    import torch
    r = gu.renderer()
    s = rt.Scene()
    w, h = 640, 360
    f = s.configure(view=5, hours=18.0, imgw=w, imgh=h, alias=True)
    full, _ = gu.render(r, s, f, want_ids=False), None
    full = full[0] if isinstance(full, tuple) else full
    comp = np.zeros((h, w), dtype=np.uint32)
    owned = np.zeros((h, w), dtype=np.int32)
    frame_dev = torch.zeros(h * w, dtype=torch.int32, device="cuda")
    for k in range(n):
        sh = rt.make_shard(n, k, tile[0], tile[1])
        part = r.render_to_host(f, shard=sh)
        again = r.render_to_host(f, shard=sh)     # second frame of the same shard: walked in cost order
        assert (again == part).all()
        yy, xx = np.mgrid[0:h, 0:w]
        mine = ((xx // tile[0]) + (yy // tile[1]) * sh.stride) % n == k
        owned += mine
        comp[mine] = part[mine]
        assert (part[~mine] == 0).all()      # a shard writes only what it owns
        # pack -> unpack reproduces the owned pixels (the gather path of bench.py --mode tiles)
        slots = r.tile_slots(sh, w, h)
        packed = torch.zeros(slots * tile[0] * tile[1], dtype=torch.int32, device="cuda")
        part_dev = gu.to_dev(part)
        r.pack_tiles(part_dev, w, h, sh, packed)
        r.unpack_tiles(packed, w, h, sh, frame_dev)
    r.synchronize()
    assert (owned == 1).all()
    assert (comp == full).all()
    assert (gu.to_host_u32(frame_dev, (h, w)) == full).all()


@pytest.mark.parametrize("fused", [1, 0])
def test_peer_frame_two_processes(fused):
    """fused = 1: rt_render_peer, the hand-shake rides in the FXAA kernel (every CTA acquires before its first peer
    store, the last CTA signals); 0: one-thread acquire / signal kernels around rt_render.
    The gather-free composite (rt_peer_frame_*): a second PROCESS opens the owner's frame through its cudaIpc
    handle and stores shard 1 of 2 straight into it while this process renders shard 0; after the flag
    hand-shake the owner's buffer equals the unsharded frame.  Three frames over two buffers, so the third
    acquire has to wait for the first release."""
    import subprocess
    import sys
    import torch
    w, h, tw, th = 640, 360, 128, 64
    r = rt.Renderer(0)
    r.set_sky(rt.synthetic_sky(256, 128))
    s = rt.Scene()
    f = s.configure(view=5, hours=18.0, imgw=w, imgh=h, alias=True)
    r.set_scene(s.objects, s.lights)
    full = torch.zeros(w * h, dtype=torch.int32, device="cuda")
    r.render(f, full)
    r.synchronize()
    got = {}
    pf = sharding.PeerFrame(r, 0, 2, w * h, lambda hb: got.setdefault("h", hb))
    child = subprocess.Popen([sys.executable, os.path.join(os.path.dirname(__file__), "peer_child.py"), got["h"].hex(),
                              str(w), str(h), str(tw), str(th), str(fused)], stdout=subprocess.PIPE, text=True)
    try:
        assert child.stdout.readline().strip() == "ready"     # its waits give up after 10 s: start together
        shard = rt.make_shard(2, 0, tw, th)
        for seq in (1, 2, 3):
            pf.render(f, shard, seq, fused=bool(fused))
            pf.wait(seq)                                       # synchronises and raises if a wait gave up
            assert torch.equal(pf.tensor(seq), full), seq
            pf.tensor(seq).zero_()
            pf.release(seq)
        assert child.wait(timeout=120) == 0
    finally:
        if child.poll() is None:
            child.kill()
        pf.close()
        r.close()


# ---------------------------------------------------------------- LBVH builder

def _check_bvh(r, n_prims):
    nodes, ids, keys = r.bvh_download()
    assert len(ids) == n_prims and sorted(ids.tolist()) == sorted(set(ids.tolist()))
    assert (np.diff(keys.astype(np.int64)) >= 0).all()
    if n_prims < 2:
        return
    assert nodes.shape == (n_prims - 1, 16)
    children = nodes[:, 12:14].copy().view(np.int32)
    seen_leaf = np.zeros(n_prims, dtype=np.int32)
    seen_node = np.zeros(n_prims - 1, dtype=np.int32)
    seen_node[0] = 1
    c = np.stack([nodes[:, 0:3], nodes[:, 6:9]], 1).astype(np.float64)      # [node][child][xyz] centre
    e = np.stack([nodes[:, 3:6], nodes[:, 9:12]], 1).astype(np.float64)     # half-extent
    assert (e >= 0).all()
    lo, hi = c - e, c + e
    for i in range(n_prims - 1):
        for k in range(2):
            ch = int(children[i, k])
            if ch < 0:
                seen_leaf[~ch] += 1
            else:
                seen_node[ch] += 1
                # a child's box (as stored in its parent) contains its own children's boxes
                tol = 1e-6 * (1.0 + np.abs(c[i, k]).max() + e[i, k].max())
                assert (lo[i, k] <= lo[ch].min(0) + tol).all() and (hi[i, k] >= hi[ch].max(0) - tol).all()
    assert (seen_leaf == 1).all() and (seen_node == 1).all()


def test_bvh_structure_builtin_scene_both_sorts():
    r = gu.renderer()
    s = rt.Scene()
    s.set_time(14.0)
    r.set_scene(s.objects[:50], s.lights)              # force a fresh build (an earlier test may have left a
    r.set_scene(s.objects, s.lights)                   # refitted tree of another time of day behind)
    assert r.bvh_info() == {"prims": 132, "nodes": 131, "planes": 1}
    _check_bvh(r, 132)
    a = r.bvh_download()
    r.set_option(_lib.RT_OPT_FORCE_RADIX, 1)           # multi-pass radix instead of the single-CTA bitonic sort
    r.set_scene(s.objects[:50], s.lights)              # topology change -> rebuild
    r.set_scene(s.objects, s.lights)
    b = r.bvh_download()
    assert (a[1] == b[1]).all() and (a[2] == b[2]).all() and (a[0][:, :14].tobytes() == b[0][:, :14].tobytes())   # raw bits: child ids are NaN patterns as floats; words 14,15 unused


@pytest.mark.parametrize("n", [1, 2, 31, 257, 4096, 4097, 100_003, 1_000_000])
def test_radix_sort_matches_library_sort(n):
    """The builder's own radix sort against torch.sort(stable=True) (a library sort; test side only)."""
    import torch
    r = gu.renderer()
    g = torch.Generator(device="cuda").manual_seed(n)
    bits = 30 if n % 2 else 6                           # many duplicates on even n: stability matters
    keys = torch.randint(0, 1 << bits, (n,), generator=g, device="cuda", dtype=torch.int32)
    vals = torch.arange(n, device="cuda", dtype=torch.int32)
    k1, v1 = keys.clone(), vals.clone()
    r.sort_pairs(k1, v1, n)
    torch.cuda.synchronize()
    ks, order = torch.sort(keys.long(), stable=True)
    assert torch.equal(k1.long(), ks) and torch.equal(v1.long(), order)


def test_refit_after_motion_equals_rebuild():
    """Sun/moon spheres and the sea level move every frame (reference scene.cpp:758-776, 709): the refit
    path must render exactly what a fresh build renders."""
    r = gu.renderer()
    s = rt.Scene()
    f = s.configure(view=6, hours=14.0, imgw=320, imgh=180, alias=True)
    gu.render(r, s, f)
    s.set_sea_level(2.0)
    f = s.configure(view=6, hours=16.5, imgw=320, imgh=180, alias=True)
    a, ida = gu.render(r, s, f)                         # same topology -> refit
    r2 = rt.Renderer(0)
    r2.set_sky(gu.sky_small())
    b, idb = gu.render(r2, s, f)                        # fresh context -> full build
    r2.close()
    assert (a == b).all() and (ida == idb).all()


def _karras_children(keys):
    """Karras 2012, one probe at a time, on the host: children of every internal node over sorted `keys` (ties broken
    by index), encoded like the device's (>= 0 node, < 0 ~leaf)."""
    n = len(keys)
    keys = [int(k) for k in keys]

    def delta(i, j):
        if j < 0 or j >= n:
            return -1
        a, b = keys[i], keys[j]
        if a == b:
            return 32 + (32 - (i ^ j).bit_length())
        return 32 - (a ^ b).bit_length()

    out = np.zeros((n - 1, 2), dtype=np.int32)
    for i in range(n - 1):
        d = 1 if delta(i, i + 1) - delta(i, i - 1) >= 0 else -1
        dmin = delta(i, i - d)
        lmax = 2
        while delta(i, i + lmax * d) > dmin:
            lmax <<= 1
        l, t = 0, lmax >> 1
        while t >= 1:
            if delta(i, i + (l + t) * d) > dmin:
                l += t
            t >>= 1
        j = i + l * d
        dnode = delta(i, j)
        s, t = 0, (l + 1) >> 1
        while True:
            if delta(i, i + (s + t) * d) > dnode:
                s += t
            if t == 1:
                break
            t = (t + 1) >> 1
        gamma = i + s * d + min(d, 0)
        lo, hi = min(i, j), max(i, j)
        out[i, 0] = ~gamma if lo == gamma else gamma
        out[i, 1] = ~(gamma + 1) if hi == gamma + 1 else gamma + 1
    return out


@pytest.mark.parametrize("cells", [24, 61])
def test_karras_topology_equals_host_karras(cells):
    """The device's Karras kernel against the algorithm restated on the host, node for node -- terrains above the PLOC
    limit (Karras is then the builder), whose Morton codes repeat (flat cells share a code at 10 bits per axis)."""
    r = gu.renderer()
    objs = rt.terrain_objects(cells)
    s = rt.Scene()
    r.set_scene(objs[:40], s.lights)                   # force a fresh build
    r.set_scene(objs, s.lights)
    n = r.bvh_info()["prims"]
    assert n > 1024
    _check_bvh(r, n)
    nodes, ids, keys = r.bvh_download()
    got = nodes[:, 12:14].copy().view(np.int32)
    assert (got == _karras_children(keys)).all()


def test_terrain_bvh_equals_bruteforce():
    """Config 4's generator at a size the brute-force kernel finishes quickly: 2*96*96+3 objects."""
    r = gu.renderer()
    objs = rt.terrain_objects(96)
    s = rt.Scene()
    s.set_pose((324.4, 330.0, -84.0), 141.2, 30.0)   # positive verAngle looks DOWN (reference scene.cpp:114)
    s.set_time(14.0)
    f = s.frame(320, 180)
    f.alias = 1
    r.set_scene(objs, s.lights)
    _check_bvh(r, len(objs) - 1)
    a, ida = r.render_to_host(f, want_ids=True)
    r.set_option(_lib.RT_OPT_TRAVERSAL, 1)
    b, idb = r.render_to_host(f, want_ids=True)
    assert (ida == idb).all() and (a == b).all()
    assert (ida > 0).mean() > 0.5                        # the terrain fills most of the view
    o = ou.oracle_render(objs, s.lights, f, gu.sky_small(), x_step=4, y_step=4, want_out=False)
    m = o["ids"][::4, ::4] != ida[::4, ::4]
    assert m.mean() <= 1e-3


def _grid_on_off(r, f, objs, lights):
    r.set_scene(objs, lights)
    r.set_option(_lib.RT_OPT_SHADOW_BINS, 2)       # build the grid at once (0 waits for a second unchanged frame)
    a = r.render_to_host(f, want_ids=False).copy()
    r.set_option(_lib.RT_OPT_SHADOW_BINS, 1)
    b = r.render_to_host(f, want_ids=False).copy()
    r.set_option(_lib.RT_OPT_SHADOW_BINS, 0)
    return a, b


def test_light_grid_of_big_scenes_changes_tested_primitives_only():
    """Shadow rays of the HBM tracer look their light's direction bin up in a cube-map grid (light_grid.cu: CSR lists
    built by a quadtree descent per primitive with the cull geometry of the small scenes' light bins) instead of
    traversing the LBVH (RT_OPT_SHADOW_BINS 0).  Conservative culling: every frame must equal the traversal's, and
    the brute-force kernel's, bit for bit -- the terrain at several times of day (sun high, grazing, below the
    horizon), lights moved into and under the terrain and next to triangles (a primitive spread over too many bins
    switches that light's grid off), the built-in scene forced through the HBM tracer with random lights, three
    lights (the third has no grid), a changed scene with unchanged lights."""
    r = gu.renderer()
    s = rt.Scene()
    s.set_pose((324.4, 330.0, -84.0), 141.2, 30.0)
    objs = rt.terrain_objects(128)
    for hours in (14.0, 7.0, 18.5, 1.0, 5.9):
        s.set_time(hours)
        f = s.frame(320, 180)
        f.alias = 0
        a, b = _grid_on_off(r, f, objs, s.lights)
        assert (a == b).all(), (hours, int((a != b).sum()))
        if hours in (14.0, 18.5):
            r.set_option(_lib.RT_OPT_TRAVERSAL, 1)
            c = r.render_to_host(f, want_ids=False)
            r.set_option(_lib.RT_OPT_TRAVERSAL, 0)
            assert (a == c).all(), (hours, "brute force")
    rng = np.random.default_rng(5)
    s.set_time(14.0)
    f = s.frame(256, 144)
    f.alias = 0
    for k in range(10):
        lights = s.lights.copy()
        for li in range(len(lights)):
            lights[li]["pos"] = rng.uniform([-500, -50, -500], [800, 350, 700]) if k % 2 else \
                rng.uniform([-2000, 300, -2000], [2000, 2500, 2000])
        a, b = _grid_on_off(r, f, objs, lights)
        assert (a == b).all(), (k, lights["pos"], int((a != b).sum()))
    # the built-in scene through the HBM tracer: spheres, big triangles, lights anywhere
    r.set_option(_lib.RT_OPT_TRACE_IMPL, 1)
    f = s.configure(view=5, hours=14.0, imgw=320, imgh=180, alias=False)
    for k in range(12):
        lights = s.lights.copy()
        if k >= 2:
            for li in range(len(lights)):
                lights[li]["pos"] = rng.uniform([-120, -20, -120], [120, 60, 120]) if k % 2 else \
                    rng.uniform([-900, -600, -900], [900, 900, 900])
        if k == 11:
            lights = np.concatenate([lights, lights[:1]]).copy()
            lights[2]["pos"] = [40.0, 80.0, -30.0]
        a, b = _grid_on_off(r, f, s.objects, lights)
        assert (a == b).all(), (k, int((a != b).sum()))
    # same lights, other scene: the grid follows the scene
    o2 = s.objects.copy()
    o2[5]["pos"] += np.float32(3.0)
    a, b = _grid_on_off(r, f, o2, s.lights)
    assert (a == b).all()
    # default policy: the first frame after a change traverses, the second builds the grid; all three agree
    r.set_scene(s.objects, s.lights)
    frames = [r.render_to_host(f, want_ids=False).copy() for _ in range(3)]
    assert (frames[0] == frames[1]).all() and (frames[1] == frames[2]).all()
    r.set_option(_lib.RT_OPT_TRACE_IMPL, 0)


def test_persistent_packet_tracer_equals_plain_tracer():
    """The plain tracer over the 4-wide view of the LBVH (RT_OPT_WIDE_BVH 1) and the persistent tracer of big scenes (trace_big.cu: warps pulling 8x4 patches, the 32 primary rays of a patch
    walking the LBVH as one packet, per-lane stacks optionally in shared memory; RT_OPT_TRACE_IMPL 3) against the plain one-thread-per-pixel tracer (RT_OPT_TRACE_IMPL 1): pixels and hit ids identical -- the
    terrain from several poses (incl. grazing rays and a camera inside it), packets off (RT_OPT_PACKET 1), the first
    stack entries in shared memory (RT_OPT_BIG_STACK 1), packets inside the plain kernel (RT_OPT_TRACE_IMPL 4), odd frame
    sizes, screen-tile shards of several tile sizes, the built-in scene (spheres, huge triangles, mirror chains),
    degenerate scenes.  The device error word (traversal stack ran out) stays clear."""
    r = gu.renderer()
    s = rt.Scene()
    objs = rt.terrain_objects(160)

    def frames(f, o, lights, shard=None):
        outs = []
        r.set_scene(o, lights)
        # the first entry is the plain tracer over the binary hierarchy; (1, 0, 0, 1) and (4, 0, 0, 1) walk its 4-wide view
        for impl, packet, sstack, wide in ((1, 0, 0, 0), (1, 0, 0, 1), (3, 0, 0, 0), (3, 1, 0, 0), (3, 0, 1, 0), (3, 1, 1, 0),
                                           (4, 0, 0, 0), (4, 0, 0, 1)):
            r.set_option(_lib.RT_OPT_TRACE_IMPL, impl)
            r.set_option(_lib.RT_OPT_PACKET, packet)
            r.set_option(_lib.RT_OPT_BIG_STACK, sstack)
            r.set_option(_lib.RT_OPT_WIDE_BVH, wide)
            if shard is None:
                outs.append(r.render_to_host(f, want_ids=True))
            else:
                img = r.render_to_host(f, shard=shard, out=np.zeros(f.imgw * f.imgh, dtype=np.uint32)).copy()
                outs.append((img, img))
        r.set_option(_lib.RT_OPT_TRACE_IMPL, 0)
        r.set_option(_lib.RT_OPT_PACKET, 0)
        r.set_option(_lib.RT_OPT_BIG_STACK, 0)
        r.set_option(_lib.RT_OPT_WIDE_BVH, 0)
        for k in range(1, len(outs)):
            assert (outs[0][1] == outs[k][1]).all(), (k, int((outs[0][1] != outs[k][1]).sum()))
            assert (outs[0][0] == outs[k][0]).all(), (k, int((outs[0][0] != outs[k][0]).sum()))
        return outs[0]

    rng = np.random.default_rng(17)
    poses = [((324.4, 330.0, -84.0), 141.2, 35.0), ((324.4, 330.0, -84.0), 141.2, 2.0), ((0.0, 20.0, 0.0), 10.0, 0.0),
             ((-400.0, 500.0, 600.0), 300.0, 44.0), ((100.0, -5.0, 100.0), 45.0, -20.0)]
    for k, (pos, hor, ver) in enumerate(poses):
        s.set_pose(pos, hor, ver)
        s.set_time(float(rng.uniform(0, 24)))
        for (w, h, alias) in ((480, 270, False), (333, 187, True)):
            s.set_antialiasing(alias)
            img, ids = frames(s.frame(w, h), objs, s.lights)
        assert k != 0 or (ids >= 1).mean() > 0.5
    # the auto choice for a big scene (the plain kernel) and the persistent tracer agree
    s.set_pose(*poses[0])
    s.set_antialiasing(False)
    f = s.frame(480, 270)
    r.set_scene(objs, s.lights)
    a = r.render_to_host(f).copy()
    r.set_option(_lib.RT_OPT_TRACE_IMPL, 3)
    assert (a == r.render_to_host(f)).all()
    r.set_option(_lib.RT_OPT_TRACE_IMPL, 0)
    # shards: tiles large enough for the four-corner shortcut, and tiles too small for it
    s.set_antialiasing(True)
    f = s.frame(640, 360)
    for n, sid, tw, th in ((2, 1, 64, 32), (4, 3, 128, 64), (8, 5, 64, 32), (3, 2, 8, 4), (2, 0, 2, 2)):
        frames(f, objs, s.lights, shard=rt.make_shard(n, sid, tw, th))
    # the built-in scene through it, both views, and a few degenerate scenes
    for view in (5, 6):
        for hours in (14.0, 1.0, 9.0):
            frames(s.configure(view=view, hours=hours, imgw=640, imgh=360, alias=False), s.objects, s.lights)
    f = s.configure(view=5, hours=14.0, imgw=160, imgh=90, alias=False)
    base = s.objects.copy()
    for subset in ([0], [0, 129], [129], [5], [0, 131, 132], list(range(11, 33))):
        frames(f, base[subset], s.lights)
    assert r.device_errors() == 0
