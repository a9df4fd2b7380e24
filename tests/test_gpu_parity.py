"""GPU parity of the CUDA hot path (through the C-ABI) against the oracle and the reference goldens.  (Against the
reference's own kernel rebuilt for sm_100: tests/test_gpu_reference_kernel.py.)

Gates (BASELINE.json north_star): primary hit ids equal except near-tie pixels; 8-bit RGB within 1 LSB
on >= 99.9 % of pixels and PSNR >= 50 dB."""
import ctypes as C
import os

import numpy as np
import pytest

import gpu_util as gu
import oracle_util as ou
import raytracing_cuda_b200 as rt
from raytracing_cuda_b200 import _lib, sharding
from raytracing_cuda_b200.structs import Camera, Float3, Light, Object

pytestmark = pytest.mark.gpu

ID_MISMATCH_MAX = 2e-4       # fraction of pixels whose primary hit id may differ (near ties at shared edges)
WITHIN_1LSB_MIN = 0.999
PSNR_MIN = 50.0


def _gate(img, ref, ids=None, ref_ids=None, what="", within=WITHIN_1LSB_MIN):
    m = ou.compare_images(img, ref)
    assert m["frac_within_1lsb"] >= within, (what, m)
    assert m["psnr_db"] >= PSNR_MIN, (what, m)
    if ids is not None:
        bad = float((ids != ref_ids).mean())
        assert bad <= ID_MISMATCH_MAX, (what, "hit-id mismatch fraction", bad)
    return m


@pytest.mark.parametrize("view", [5, 6])
@pytest.mark.parametrize("tod", list(rt.TIME_PRESETS))
def test_bvh_equals_bruteforce_bit_exact(view, tod):
    """The LBVH may only change WHICH primitives are tested: ids and pixels identical to the brute-force kernel."""
    r = gu.renderer()
    s = rt.Scene()
    f = s.configure(view=view, hours=rt.TIME_PRESETS[tod], imgw=640, imgh=360, alias=True)
    img, ids = gu.render(r, s, f)
    r.set_option(_lib.RT_OPT_TRAVERSAL, 1)
    img_b, ids_b = gu.render(r, s, f)
    assert (ids == ids_b).all()
    assert (img == img_b).all()


@pytest.mark.parametrize("view", [5, 6])
@pytest.mark.parametrize("tod", list(rt.TIME_PRESETS))
def test_gpu_vs_reference_golden_frames(view, tod):
    FR = np.load(os.path.join(ou.GOLDEN, "frames.npz"))
    r = gu.renderer()
    s = rt.Scene()
    f = s.configure(view=view, hours=rt.TIME_PRESETS[tod], imgw=320, imgh=180, alias=True)
    img, ids = gu.render(r, s, f)
    key = "v%d_%s" % (view, tod)
    # goldens: the reference's kernel.cu text compiled for the HOST (IEEE divides, libm) at 320x180; the north_star
    # gates apply unrelaxed (the comparison with the reference's own GPU kernel is tests/test_gpu_reference_kernel.py)
    _gate(img, FR[key + "_out"], ids, FR[key + "_ids"].astype(np.int32), key)
    f.alias = 0
    img0, _ = gu.render(r, s, f)
    _gate(img0, FR[key + "_base"], what=key + " base")


@pytest.mark.parametrize("cfg", [(6, 14.0, 1280, 720, False, -4.5), (5, 18.0, 640, 360, True, -4.5),
                                 (6, 14.0, 480, 270, True, 7.5), (5, 9.0, 320, 180, True, -4.5),
                                 (5, 21.3, 333, 187, True, 1.5)])
def test_gpu_vs_oracle(cfg):
    """Config 1 of BASELINE.json (Mountains 1280x720 AA off) and reduced cases incl. a cross-fade hour,
    raised sea levels and an odd size (width % 4 != 0 -> shared-memory FXAA)."""
    view, hours, w, h, alias, sea = cfg
    r = gu.renderer()
    s = rt.Scene()
    s.set_sea_level(sea)
    f = s.configure(view=view, hours=hours, imgw=w, imgh=h, alias=alias)
    img, ids = gu.render(r, s, f)
    o = ou.oracle_render(s.objects, s.lights, f, gu.sky_small())
    _gate(img, o["out"], ids, o["ids"], str(cfg))


def test_counters_match_oracle_ray_counts():
    r = gu.renderer()
    r.set_option(_lib.RT_OPT_COUNTERS, 1)
    s = rt.Scene()
    f = s.configure(view=5, hours=14.0, imgw=320, imgh=180, alias=False)
    gu.render(r, s, f)
    c = r.counters()
    o = ou.oracle_render(s.objects, s.lights, f, gu.sky_small())["counters"]
    assert c["rays_primary"] == o["rays_primary"] == 320 * 180
    assert abs(c["rays_shadow"] - o["rays_shadow"]) <= 0.002 * o["rays_shadow"]
    assert abs(c["rays_mirror"] - o["rays_mirror"]) <= 0.002 * o["rays_mirror"]
    assert abs(c["sky_lookups"] - o["sky_lookups"]) <= 0.002 * o["sky_lookups"]
    # the LBVH does far fewer primitive tests than the reference's linear loops
    assert c["tri_tests"] + c["sphere_tests"] + c["plane_tests"] < 0.25 * o["prim_tests"]


def test_sky_all4_is_bit_identical():
    r = gu.renderer()
    s = rt.Scene()
    for hours in (14.0, 9.0, 5.0):
        f = s.configure(view=6, hours=hours, imgw=320, imgh=180, alias=False)
        a, _ = gu.render(r, s, f)
        r.set_option(_lib.RT_OPT_SKY_ALL4, 1)
        b, _ = gu.render(r, s, f)
        r.set_option(_lib.RT_OPT_SKY_ALL4, 0)
        assert (a == b).all()


def test_generic_scene_edge_cases():
    """Empty BVH (planes only), a single primitive, origin inside a sphere, emissive-only."""
    r = gu.renderer()
    s = rt.Scene()
    f = s.configure(view=5, hours=14.0, imgw=160, imgh=90, alias=False)
    objs = s.objects.copy()
    for subset in ([0], [0, 129], [129], [5], [0, 131, 132], list(range(11, 33))):
        o = objs[subset]
        img, ids = gu.render(r, s, f, objects=o)
        ref = ou.oracle_render(o, s.lights, f, gu.sky_small())
        _gate(img, ref["out"], ids, ref["ids"], str(subset))
    # camera inside the igloo sphere (index 129): negative-t near root must still win (kernel.cu:62)
    s.set_pose((4.0, -3.0, -4.0), 30.0, 5.0)
    f = s.frame(160, 90)
    f.alias = 0
    img, ids = gu.render(r, s, f)
    ref = ou.oracle_render(s.objects, s.lights, f, gu.sky_small())
    _gate(img, ref["out"], ids, ref["ids"], "inside igloo")
    assert (ids == 129).mean() > 0.3


def _both_impls(r, s, f, shard=None, objects=None):
    r.set_option(_lib.RT_OPT_TRACE_IMPL, 0)
    a = gu.render(r, s, f, shard=shard, objects=objects)
    r.set_option(_lib.RT_OPT_TRACE_IMPL, 1)
    b = gu.render(r, s, f, shard=shard, objects=objects)
    r.set_option(_lib.RT_OPT_TRACE_IMPL, 0)
    return a, b


@pytest.mark.parametrize("cfg", [(5, 14.0, 1920, 1080, -4.5), (5, 1.0, 1280, 720, -4.5), (6, 18.0, 1280, 720, 7.5),
                                 (6, 6.0, 641, 359, -4.5), (5, 9.0, 33, 9, -4.5), (6, 21.0, 2, 2, 3.5)])
def test_shared_memory_tracer_equals_hbm_tracer(cfg):
    """The persistent shared-memory tracer (tile-frustum candidate lists for primary rays, shared-memory
    LBVH for the bounces; auto-selected for small scenes) must agree with the HBM-resident LBVH tracer on
    every pixel and every hit id: both may only change WHICH primitives get tested."""
    view, hours, w, h, sea = cfg
    r = gu.renderer()
    s = rt.Scene()
    s.set_sea_level(sea)
    f = s.configure(view=view, hours=hours, imgw=w, imgh=h, alias=False)
    (img_a, ids_a), (img_b, ids_b) = _both_impls(r, s, f)
    assert (ids_a == ids_b).all(), int((ids_a != ids_b).sum())
    assert (img_a == img_b).all(), int((img_a != img_b).sum())


def test_shared_memory_tracer_poses_shards_and_subsets():
    r = gu.renderer()
    s = rt.Scene()
    rng = np.random.default_rng(7)
    # random camera poses all over (and outside) the scene, including straight up / down
    for k in range(24):
        pos = tuple(float(v) for v in rng.uniform([-700, -3, -700], [900, 400, 800]))
        s.set_pose(pos, float(rng.uniform(0, 360)), float(rng.uniform(-44, 44)))
        s.set_time(float(rng.uniform(0, 24)))
        f = s.frame(320, 180)
        f.alias = 0
        (img_a, ids_a), (img_b, ids_b) = _both_impls(r, s, f)
        assert (ids_a == ids_b).all(), (k, pos, int((ids_a != ids_b).sum()))
        assert (img_a == img_b).all(), (k, pos)
    # screen-tile shards
    s = rt.Scene()
    f = s.configure(view=5, hours=14.0, imgw=640, imgh=360, alias=True)
    for n, sid in ((2, 1), (4, 3), (8, 5)):
        sh = rt.make_shard(n, sid, 64, 32)
        (img_a, ids_a), (img_b, ids_b) = _both_impls(r, s, f, shard=sh)
        own = sharding.owner_map(640, 360, 64, 32, n, sh.stride) == sid
        assert (img_a[own] == img_b[own]).all()
    # the deep-hierarchy fallback of the shared-memory tracer (linear loop for the bounces)
    f2 = s.configure(view=5, hours=18.0, imgw=320, imgh=180, alias=False)
    r.set_option(_lib.RT_OPT_TRACE_IMPL, 2)
    img_c, ids_c = gu.render(r, s, f2)
    r.set_option(_lib.RT_OPT_TRACE_IMPL, 1)
    img_d, ids_d = gu.render(r, s, f2)
    r.set_option(_lib.RT_OPT_TRACE_IMPL, 0)
    assert (ids_c == ids_d).all() and (img_c == img_d).all()
    # subsets: planes only, one primitive, spheres only
    objs = s.objects.copy()
    f.alias = 0
    for subset in ([0], [0, 129], [129], [5], [0, 131, 132], list(range(11, 33)), list(range(0, 133, 2))):
        (img_a, ids_a), (img_b, ids_b) = _both_impls(r, s, f, objects=objs[subset])
        assert (ids_a == ids_b).all(), subset
        assert (img_a == img_b).all(), subset


def _bins_on_off(r, s, f, objects=None, lights=None):
    r.set_option(_lib.RT_OPT_SHADOW_BINS, 0)
    a = gu.render(r, s, f, objects=objects, lights=lights)
    r.set_option(_lib.RT_OPT_SHADOW_BINS, 1)
    b = gu.render(r, s, f, objects=objects, lights=lights)
    r.set_option(_lib.RT_OPT_SHADOW_BINS, 0)
    return a, b


def test_light_bins_change_tested_primitives_only():
    """Shadow rays of the shared-memory tracer look up per-light direction bins (a cube map around each light
    with the primitives a line of that direction can hit) instead of traversing the hierarchy
    (RT_OPT_SHADOW_BINS 0).  Conservative culling: every frame must equal the traversal's (and, in the tests
    above, the HBM tracer's and the oracle's) bit for bit -- all hours of both views, lights moved into and
    under the scene, random clouds of spheres and triangles around the lights, more lights than have bins."""
    r = gu.renderer()
    s = rt.Scene()
    for view in (5, 6):
        for hours in np.arange(0.0, 24.0, 0.75):
            f = s.configure(view=view, hours=float(hours), imgw=480, imgh=270, alias=False)
            (img_a, ids_a), (img_b, ids_b) = _bins_on_off(r, s, f)
            assert (ids_a == ids_b).all() and (img_a == img_b).all(), (view, hours, int((img_a != img_b).sum()))
    # lights anywhere: inside the island, under the sea, next to triangles, far away
    rng = np.random.default_rng(11)
    f = s.configure(view=5, hours=14.0, imgw=320, imgh=180, alias=False)
    for k in range(16):
        lights = s.lights.copy()
        for li in range(len(lights)):
            lights[li]["pos"] = rng.uniform([-120, -20, -120], [120, 60, 120]) if k % 2 else \
                rng.uniform([-900, -600, -900], [900, 900, 900])
        (img_a, ids_a), (img_b, ids_b) = _bins_on_off(r, s, f, lights=lights)
        assert (img_a == img_b).all(), (k, lights["pos"], int((img_a != img_b).sum()))
        r.set_option(_lib.RT_OPT_TRACE_IMPL, 1)
        img_c, _ = gu.render(r, s, f, lights=lights)
        r.set_option(_lib.RT_OPT_TRACE_IMPL, 0)
        assert (img_a == img_c).all(), (k, "hbm")
    # random clouds: many small primitives stacked along the light directions (bins overflow -> traversal)
    base = s.objects.copy()
    for k in range(8):
        n = int(rng.integers(20, 160))
        o = np.zeros(n + 1, dtype=base.dtype)
        o[0] = base[0]                                  # the sea plane
        src_t = base[base["type"] == 2][0]
        src_s = base[(base["type"] == 0) & (base["light"] == 0)][0]
        for i in range(1, n + 1):
            c = rng.uniform([-90, -2, 20], [-20, 40, 110])
            if rng.random() < 0.5:
                o[i] = src_s
                o[i]["pos"] = c
                o[i]["size"] = [rng.uniform(0.3, 6.0)] * 3
            else:
                o[i] = src_t
                a, b = rng.normal(size=3) * rng.uniform(1, 12), rng.normal(size=3) * rng.uniform(1, 12)
                o[i]["pos"], o[i]["size"], o[i]["third"] = c, c + a, c + b
            o[i]["mirror"] = float(rng.choice([0.0, 0.3]))
        (img_a, ids_a), (img_b, ids_b) = _bins_on_off(r, s, f, objects=o)
        assert (ids_a == ids_b).all() and (img_a == img_b).all(), (k, n, int((img_a != img_b).sum()))
    # five lights: the fifth has no bins and traverses
    lights = np.concatenate([s.lights, s.lights, s.lights[:1]]).copy()
    for li in range(5):
        lights[li]["pos"] = rng.uniform([-300, 20, -300], [300, 500, 300])
        lights[li]["intensity"] = 0.3
    (img_a, _), (img_b, _) = _bins_on_off(r, s, f, lights=lights)
    assert (img_a == img_b).all()


def _mirror_on_off(r, s, f, objects=None, lights=None):
    r.set_option(_lib.RT_OPT_MIRROR_BINS, 0)
    a = gu.render(r, s, f, objects=objects, lights=lights)
    r.set_option(_lib.RT_OPT_MIRROR_BINS, 1)
    b = gu.render(r, s, f, objects=objects, lights=lights)
    r.set_option(_lib.RT_OPT_MIRROR_BINS, 0)
    return a, b


def test_mirror_bins_change_tested_primitives_only():
    """The ray a primary ray sends off plane 0 (the sea, reference scene.cpp:183-190; bounce at kernel.cu:209-215)
    tests the candidate list of its screen bin's MIRRORED frustum (apex = the eye's image behind the plane)
    instead of traversing the hierarchy (RT_OPT_MIRROR_BINS 0).  Conservative culling: every frame must equal
    the traversal's bit for bit -- all hours of both views, the sea-level sweep of BASELINE config 3 (including
    levels above the camera), cameras close to / under the plane, tilted unit and non-unit plane normals (the
    latter switch the lists off), random clouds above and below the plane, no plane at all."""
    r = gu.renderer()
    s = rt.Scene()
    for view in (5, 6):
        for hours in np.arange(0.0, 24.0, 1.5):
            f = s.configure(view=view, hours=float(hours), imgw=480, imgh=270, alias=False)
            (img_a, ids_a), (img_b, ids_b) = _mirror_on_off(r, s, f)
            assert (ids_a == ids_b).all() and (img_a == img_b).all(), (view, hours, int((img_a != img_b).sum()))
    for sea in (-4.5, -2.5, -0.5, 1.5, 3.5, 5.5, 7.5, 9.5, 11.5, 12.40, 12.41, 12.42, 14.0, 40.0):
        f = s.configure(view=6, hours=14.0, imgw=640, imgh=360, sea_level=sea, alias=False)
        (img_a, ids_a), (img_b, ids_b) = _mirror_on_off(r, s, f)
        assert (ids_a == ids_b).all() and (img_a == img_b).all(), (sea, int((img_a != img_b).sum()))
    # against the HBM tracer too (it has no lists at all)
    f = s.configure(view=5, hours=18.0, imgw=1920, imgh=1080, alias=False)
    img_a, ids_a = gu.render(r, s, f)
    r.set_option(_lib.RT_OPT_TRACE_IMPL, 1)
    img_c, ids_c = gu.render(r, s, f)
    r.set_option(_lib.RT_OPT_TRACE_IMPL, 0)
    assert (ids_a == ids_c).all() and (img_a == img_c).all()
    # moving camera: over, on and under the sea, looking up and down
    rng = np.random.default_rng(23)
    s.configure(view=5, hours=10.0, imgw=320, imgh=180, alias=False)
    for k in range(24):
        pos = rng.uniform([-300, -30, -300], [400, 60, 300])
        if k % 4 == 0:
            pos[1] = -4.5 + rng.uniform(-0.02, 0.02)
        s.set_pose(tuple(float(v) for v in pos), float(rng.uniform(0, 360)), float(rng.uniform(-44, 44)))
        f = s.frame(320, 180)
        (img_a, ids_a), (img_b, ids_b) = _mirror_on_off(r, s, f)
        assert (ids_a == ids_b).all() and (img_a == img_b).all(), (k, pos, int((img_a != img_b).sum()))
    # tilted planes: unit normals keep the lists, non-unit normals (not a reflection any more) switch them off
    f = s.configure(view=5, hours=14.0, imgw=320, imgh=180, alias=False)
    base = s.objects.copy()
    for k in range(12):
        o = base.copy()
        nrm = np.array([rng.uniform(-0.4, 0.4), 1.0, rng.uniform(-0.4, 0.4)])
        nrm = nrm / np.linalg.norm(nrm)
        if k % 3 == 1:
            nrm = nrm * rng.uniform(0.5, 2.0)
        if k % 3 == 2:
            nrm = -nrm
        o[0]["size"] = nrm.astype(np.float32)
        o[0]["pos"] = [rng.uniform(-50, 50), rng.uniform(-8, 3), rng.uniform(-50, 50)]
        (img_a, ids_a), (img_b, ids_b) = _mirror_on_off(r, s, f, objects=o)
        assert (ids_a == ids_b).all() and (img_a == img_b).all(), (k, nrm, int((img_a != img_b).sum()))
    # random clouds around the plane (lists overflow -> traversal), and a scene without any plane
    for k in range(8):
        n = int(rng.integers(10, 200))
        o = np.zeros(n + 1, dtype=base.dtype)
        o[0] = base[0]
        src_t = base[base["type"] == 2][0]
        src_s = base[(base["type"] == 0) & (base["light"] == 0)][0]
        for i in range(1, n + 1):
            c = rng.uniform([-110, -12, 10], [0, 30, 130])
            if rng.random() < 0.4:
                o[i] = src_s
                o[i]["pos"] = c
                o[i]["size"] = [rng.uniform(0.3, 5.0)] * 3
            else:
                o[i] = src_t
                a, b = rng.normal(size=3) * rng.uniform(1, 15), rng.normal(size=3) * rng.uniform(1, 15)
                o[i]["pos"], o[i]["size"], o[i]["third"] = c, c + a, c + b
            o[i]["mirror"] = float(rng.choice([0.0, 0.4]))
        if k == 7:
            o = o[1:]
        (img_a, ids_a), (img_b, ids_b) = _mirror_on_off(r, s, f, objects=o)
        assert (ids_a == ids_b).all() and (img_a == img_b).all(), (k, n, int((img_a != img_b).sum()))
        r.set_option(_lib.RT_OPT_TRACE_IMPL, 1)
        img_c, _ = gu.render(r, s, f, objects=o)
        r.set_option(_lib.RT_OPT_TRACE_IMPL, 0)
        assert (img_a == img_c).all(), (k, "hbm")
    # a sharded frame uses the same lists
    f = s.configure(view=6, hours=14.0, imgw=640, imgh=360, alias=True)
    r.set_scene(s.objects, s.lights)
    sh = rt.make_shard(2, 1, 64, 32)
    img_a = r.render_to_host(f, shard=sh, out=np.zeros(640 * 360, dtype=np.uint32)).copy()
    r.set_option(_lib.RT_OPT_MIRROR_BINS, 1)
    img_b = r.render_to_host(f, shard=sh, out=np.zeros(640 * 360, dtype=np.uint32)).copy()
    r.set_option(_lib.RT_OPT_MIRROR_BINS, 0)
    assert (img_a == img_b).all() and img_a.any()


def test_ray_compaction_changes_scheduling_only():
    """Rays of the second and later generations that have to traverse the hierarchy are parked in a per-CTA
    shared-memory queue and traced 32 at a time by whichever warp finds a full batch (RT_OPT_COMPACTION 1).  Which
    thread finishes a pixel must not matter: pixels, hit ids and the ray / test counters equal the lane-per-pixel
    walk (RT_OPT_COMPACTION 0, the default) -- both views round the clock, the reflection-heavy sea levels of config 3, tiny and
    odd frame sizes (partial batches only), shards, the deep-hierarchy fallback, mirror / light bins off (every bounce
    and shadow ray traverses: the queue overflows and lanes finish rays themselves), random mirror clouds."""
    r = gu.renderer()
    s = rt.Scene()

    def both(f, **kw):
        r.set_option(_lib.RT_OPT_COMPACTION, 0)
        a = gu.render(r, s, f, **kw)
        r.set_option(_lib.RT_OPT_COMPACTION, 1)
        b = gu.render(r, s, f, **kw)
        r.set_option(_lib.RT_OPT_COMPACTION, 0)
        assert (a[1] == b[1]).all(), int((a[1] != b[1]).sum())
        assert (a[0] == b[0]).all(), int((a[0] != b[0]).sum())
        return a

    for view in (5, 6):
        for hours in np.arange(0.0, 24.0, 3.0):
            both(s.configure(view=view, hours=float(hours), imgw=960, imgh=540, alias=False))
    for sea in (-4.5, 1.5, 7.5, 11.5):
        both(s.configure(view=6, hours=14.0, imgw=1280, imgh=720, sea_level=sea, alias=True))
    both(s.configure(view=5, hours=14.0, imgw=1920, imgh=1080, sea_level=-4.5, alias=True))
    for (w, h) in ((33, 9), (2, 2), (641, 359), (8, 4), (64, 36)):
        both(s.configure(view=5, hours=18.0, imgw=w, imgh=h, alias=False))
    f = s.configure(view=5, hours=14.0, imgw=640, imgh=360, alias=True)
    for n, sid in ((2, 1), (4, 3)):
        sh = rt.make_shard(n, sid, 64, 32)
        r.set_scene(s.objects, s.lights)
        outs = []
        for c in (0, 1):
            r.set_option(_lib.RT_OPT_COMPACTION, c)
            outs.append(r.render_to_host(f, shard=sh, out=np.zeros(640 * 360, dtype=np.uint32)).copy())
        r.set_option(_lib.RT_OPT_COMPACTION, 0)
        assert (outs[0] == outs[1]).all() and outs[0].any()
    # every secondary ray traverses: bins off, then the deep-hierarchy fallback (linear loops) on top
    f = s.configure(view=5, hours=9.0, imgw=800, imgh=450, alias=False)
    r.set_option(_lib.RT_OPT_MIRROR_BINS, 1)
    r.set_option(_lib.RT_OPT_SHADOW_BINS, 1)
    both(f)
    r.set_option(_lib.RT_OPT_TRACE_IMPL, 2)
    both(f)
    for k in (_lib.RT_OPT_TRACE_IMPL, _lib.RT_OPT_MIRROR_BINS, _lib.RT_OPT_SHADOW_BINS):
        r.set_option(k, 0)
    # counters: the same rays and the same tests, whoever runs them
    f = s.configure(view=5, hours=14.0, imgw=960, imgh=540, alias=False)
    cs = []
    for c in (0, 1):
        r.set_option(_lib.RT_OPT_COMPACTION, c)
        r.set_option(_lib.RT_OPT_COUNTERS, 1)
        gu.render(r, s, f)
        cs.append(r.counters())
        r.set_option(_lib.RT_OPT_COUNTERS, 0)
    r.set_option(_lib.RT_OPT_COMPACTION, 0)
    for k in ("rays_primary", "rays_shadow", "rays_mirror", "sky_lookups", "shadow_skipped", "node_tests", "plane_tests"):
        assert cs[0][k] == cs[1][k], (k, cs[0], cs[1])
    # leaf tests of shadow rays stop at the first occluder of their light bin, and the pre-pass fills a bin in whatever
    # order its atomics land (rebuilt between the two frames): some tests more or less (2e-4 of them has been seen),
    # never other pixels
    for k in ("tri_tests", "sphere_tests"):
        assert abs(cs[0][k] - cs[1][k]) <= 1e-3 * cs[0][k], (k, cs[0], cs[1])
    # random clouds of mirrors: long chains everywhere
    rng = np.random.default_rng(31)
    base = s.objects.copy()
    f = s.configure(view=5, hours=14.0, imgw=480, imgh=270, alias=False)
    for k in range(6):
        n = int(rng.integers(30, 200))
        o = np.zeros(n + 1, dtype=base.dtype)
        o[0] = base[0]
        src_t = base[base["type"] == 2][0]
        src_s = base[(base["type"] == 0) & (base["light"] == 0)][0]
        for i in range(1, n + 1):
            c = rng.uniform([-110, -6, 10], [0, 30, 130])
            if rng.random() < 0.4:
                o[i] = src_s
                o[i]["pos"] = c
                o[i]["size"] = [rng.uniform(0.5, 6.0)] * 3
            else:
                o[i] = src_t
                a, b = rng.normal(size=3) * rng.uniform(2, 15), rng.normal(size=3) * rng.uniform(2, 15)
                o[i]["pos"], o[i]["size"], o[i]["third"] = c, c + a, c + b
            o[i]["mirror"] = float(rng.choice([0.0, 0.3, 0.8]))
        both(f, objects=o)


def test_order_history_changes_scheduling_only():
    """Whole frames of the shared-memory tracer are walked most-expensive-patch-first using the costs the
    previous frame of the same size measured (RT_OPT_ORDER 0).  The frames must not depend on it: same pixels
    and ids as the plain bottom-up walk, across repeated frames, a moving camera, changing sizes and an
    interleaved sharded frame."""
    r = gu.renderer()
    s = rt.Scene()
    for (view, hours, w, h) in ((5, 14.0, 1920, 1080), (6, 18.0, 641, 359), (5, 1.0, 33, 9)):
        f = s.configure(view=view, hours=hours, imgw=w, imgh=h, alias=True)
        r.set_option(_lib.RT_OPT_ORDER, 1)
        img_ref, ids_ref = gu.render(r, s, f)
        r.set_option(_lib.RT_OPT_ORDER, 0)
        for k in range(4):
            if k == 2:   # a sharded frame and a frame of another size in between
                gu.render(r, s, f, shard=rt.make_shard(2, 1, 64, 32))
                gu.render(r, s, s.configure(view=view, hours=hours, imgw=160, imgh=90, alias=True))
                f = s.configure(view=view, hours=hours, imgw=w, imgh=h, alias=True)
            img, ids = gu.render(r, s, f)
            assert (ids == ids_ref).all(), (w, h, k)
            assert (img == img_ref).all(), (w, h, k)
    # the history follows a moving camera
    for k in range(6):
        s.set_pose((-56.0 + 20 * k, 2.2 + 3 * k, 72.0 - 10 * k), 309.0 - 15 * k, -7.0 + 2 * k)
        f = s.frame(640, 360)
        r.set_option(_lib.RT_OPT_ORDER, 0)
        img, ids = gu.render(r, s, f)
        r.set_option(_lib.RT_OPT_ORDER, 1)
        img_ref, ids_ref = gu.render(r, s, f)
        r.set_option(_lib.RT_OPT_ORDER, 0)
        assert (ids == ids_ref).all() and (img == img_ref).all(), k


@pytest.mark.parametrize("alias", [True, False])
def test_back_to_back_frames_overlap_safely(alias):
    """Consecutive frames of one context with an unchanged scene: the next frame's binning pre-pass is launched
    programmatically dependent and runs while the previous frame's FXAA / copy kernel finishes; the tracer must
    still never overwrite the base image under it.  40 frames alternating between two cameras, no host
    synchronisation in between, every output compared with the frame rendered alone."""
    import torch
    r = gu.renderer()
    s = rt.Scene()
    w, h = 1280, 720
    frames, refs = [], []
    for view in (5, 6):
        f = s.configure(view=view, hours=14.0, imgw=w, imgh=h, alias=alias)
        fr = type(f)()
        C.memmove(C.byref(fr), C.byref(f), C.sizeof(f))
        frames.append(fr)
    r.set_scene(s.objects, s.lights)
    for fr in frames:
        o = torch.zeros(w * h, dtype=torch.int32, device="cuda")
        r.render(fr, o)
        r.synchronize()
        refs.append(o)
    outs = [torch.zeros(w * h, dtype=torch.int32, device="cuda") for _ in range(40)]
    for k, o in enumerate(outs):
        r.render(frames[k & 1], o)
    r.synchronize()
    for k, o in enumerate(outs):
        assert torch.equal(o, refs[k & 1]), k


# ---------------------------------------------------------------- legacy boundary

def _legacy_call(lib, out_ptr, s, f, sky):
    objs = s.objects
    lights = s.lights
    cam = Camera()
    C.memmove(C.byref(cam), C.byref(f.cam), C.sizeof(Camera))
    amb = Float3(*f.ambient)
    skyv = (C.c_float * 4)(*f.skyVars)
    lib.launchKernel(out_ptr, f.imgw, f.imgh, cam, C.cast(objs.ctypes.data, C.POINTER(Object)),
                     C.cast(lights.ctypes.data, C.POINTER(Light)), amb, skyv, sky[0].ctypes.data, sky[1].ctypes.data,
                     sky[2].ctypes.data, sky[3].ctypes.data, sky.shape[2], sky.shape[1], f.dayTime, bool(f.alias))


def test_legacy_launchKernel_matches_handle_api():
    """launchKernel(...) with the reference's exact signature (kernel.cu:406-411): device result pointer,
    host everything else, Camera and float3 by value, asynchronous."""
    import torch
    lib = _lib.load()
    r = gu.renderer()
    s = rt.Scene()
    sky = gu.sky_small()
    out = torch.zeros(640 * 360, dtype=torch.int32, device="cuda")
    for view, hours, alias in ((5, 14.0, True), (6, 18.0, False), (5, 1.0, True)):
        f = s.configure(view=view, hours=hours, imgw=640, imgh=360, alias=alias)
        _legacy_call(lib, out.data_ptr(), s, f, sky)
        torch.cuda.synchronize()
        got = out.cpu().numpy().view(np.uint32).reshape(360, 640)
        want, _ = gu.render(r, s, f)
        assert (got == want).all()


@pytest.mark.skipif(ou.ref_lib("scene") is None, reason="oracle/_ref/libref_scene.so did not travel")
def test_reference_launch_drives_the_drop_in():
    """The reference's OWN frame entry point launch() (scene.cpp:820-826), compiled from its sources,
    calling into this library's launchKernel: the drop-in claim end to end."""
    import torch
    lib = _lib.load()
    ref = ou.RefScene()
    sky = gu.sky_small()
    ref.lib.ref_scene_set_textures.argtypes = [C.c_void_p] * 4 + [C.c_int, C.c_int]
    ref.lib.ref_scene_set_textures(sky[0].ctypes.data, sky[1].ctypes.data, sky[2].ctypes.data, sky[3].ctypes.data,
                                   sky.shape[2], sky.shape[1])
    ref.lib.ref_scene_set_forward.argtypes = [C.c_void_p]
    ref.lib.ref_scene_set_forward(C.cast(lib.launchKernel, C.c_void_p))
    ref.lib.ref_scene_launch.argtypes = [C.c_void_p, C.c_int, C.c_int]
    w, h = 480, 270
    pos, hor, ver = rt.POSES[6]
    ref.lib.ref_scene_set_aspect(w, h)
    ref.lib.ref_scene_set_pose(*pos, hor, ver)
    ref.lib.ref_scene_set_time(14.0)
    ref.lib.ref_scene_set_aa(1)
    out = torch.zeros(w * h, dtype=torch.int32, device="cuda")
    ref.lib.ref_scene_launch(out.data_ptr(), w, h)          # launch() ends with cudaDeviceSynchronize
    torch.cuda.synchronize()
    got = out.cpu().numpy().view(np.uint32).reshape(h, w)
    s = rt.Scene()
    f = s.configure(view=6, hours=14.0, imgw=w, imgh=h, alias=True)
    o = ou.oracle_render(s.objects, s.lights, f, sky)
    _gate(got, o["out"], what="reference launch() -> drop-in")
