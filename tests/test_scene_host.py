"""Host scene layer (csrc/scene_host.cpp) against the reference's own scene.cpp:
golden states generated from it (tests/golden/scene_states.npz, made by tests/gen_golden.py) and,
when oracle/_ref/libref_scene.so is present, the live reference build."""
import os

import numpy as np
import pytest

import oracle_util as ou
import raytracing_cuda_b200 as rt
from raytracing_cuda_b200.structs import TRIANGLE, SPHERE, PLANE, objects_equal

G = np.load(os.path.join(ou.GOLDEN, "scene_states.npz"))


def _same_state(s, prefix):
    assert objects_equal(G[prefix + "_objects"], s.objects), prefix
    assert G[prefix + "_lights"].tobytes() == s.lights.tobytes(), prefix
    assert G[prefix + "_camera"][0].tobytes() == s.camera.tobytes(), prefix
    assert G[prefix + "_sky"].tobytes() == s.sky_vars.tobytes(), prefix
    assert G[prefix + "_ambient"].tobytes() == s.ambient.tobytes(), prefix


def test_scene_inventory():
    """133 primitives = 1 plane + 106 triangles + 26 spheres, index layout fixed."""
    s = rt.Scene()
    o = s.objects
    assert len(o) == 133
    assert (o["type"] == PLANE).sum() == 1 and (o["type"] == TRIANGLE).sum() == 106 and (o["type"] == SPHERE).sum() == 26
    assert o["type"][0] == PLANE and o["pos"][0][1] == np.float32(-4.5)
    assert (o["type"][1:11] == TRIANGLE).all() and (o["type"][11:33] == SPHERE).all()
    assert (o["type"][33:129] == TRIANGLE).all() and (o["type"][129:133] == SPHERE).all()
    assert list(np.nonzero(o["light"])[0]) == [131, 132]


def test_states_match_reference_goldens():
    s = rt.Scene()
    for case, (view, hours, sea) in enumerate(G["meta"]):
        s.set_view(int(view))
        s.set_sea_level(sea)
        s.set_time(hours)
        s.update_camera(1920, 1080)
        _same_state(s, "c%03d" % case)


def test_animate_matches_reference_goldens():
    """animate(): time advances 0.5 h/s, colours lag the sky weights by one tick (reference scene.cpp:806-816)."""
    s = rt.Scene()
    s.set_view(6)
    s.update_camera(1920, 1080)
    s.set_sea_level(-4.5)
    s.set_time(5.0)
    for tick in range(40):
        s.animate(0.75)
        p = "a%03d" % tick
        assert np.float32(s.time) == G[p + "_time"]
        assert objects_equal(G[p + "_objects"], s.objects), tick
        assert G[p + "_lights"].tobytes() == s.lights.tobytes()
        assert G[p + "_sky"].tobytes() == s.sky_vars.tobytes()
        assert G[p + "_ambient"].tobytes() == s.ambient.tobytes()
        assert G[p + "_clock"].tobytes().rstrip(b"\0") == s.clock().encode()


def test_time_presets_table():
    """sky weights and sun positions at the four presets."""
    s = rt.Scene()
    want = {6.0: (1, 0, 0, 0), 14.0: (0, 1, 0, 0), 18.0: (0, 0, 1, 0), 1.0: (0, 0, 0, 1)}
    for hours, sky in want.items():
        s.set_time(hours)
        assert tuple(s.sky_vars) == sky
    s.set_time(14.0)
    np.testing.assert_allclose(s.lights["pos"][0], (-500, 500, 500), atol=1e-3)
    np.testing.assert_allclose(s.lights["pos"][1], (-500, -500, 500), atol=1e-3)
    assert (s.objects["pos"][131] == s.lights["pos"][0]).all() and (s.objects["pos"][132] == s.lights["pos"][1]).all()


def test_controls_and_presets():
    s = rt.Scene()
    s.animate(0.0, keys=["6", "V", "O"])
    assert tuple(s.camera["pos"]) == (np.float32(324.4), np.float32(12.41), np.float32(-84.0))
    s.animate(1.0, keys=["UP"])           # sea rises 2 units/s (reference scene.cpp:25, 709)
    assert s.objects["pos"][0][1] == np.float32(-2.5)
    t0 = s.time
    s.animate(1.0)                        # paused by 'O'
    assert s.time == t0
    s.animate(1.0, keys=["RIGHT"])        # 0.5 * 4 h/s
    assert abs(s.time - (t0 + 2.0)) < 1e-5
    s.animate(0.5, keys=["W", "SHIFT"])   # 50 units/s, doubled
    assert s.frame(64, 36).alias == 0


@pytest.mark.skipif(ou.ref_lib("scene") is None, reason="oracle/_ref/libref_scene.so not built (needs /root/reference)")
def test_live_reference_scene_random_states():
    ref = ou.RefScene()
    s = rt.Scene()
    rng = np.random.default_rng(3)
    for _ in range(200):
        hours = float(np.float32(rng.uniform(0, 24)))
        pos = [float(np.float32(v)) for v in rng.uniform(-300, 300, 3)]
        hor, ver = float(np.float32(rng.uniform(0, 360))), float(np.float32(rng.uniform(-44, 44)))
        w, h = int(rng.integers(16, 4000)), int(rng.integers(16, 2200))
        ref.lib.ref_scene_set_aspect(w, h)
        ref.lib.ref_scene_set_pose(*pos, hor, ver)
        ref.lib.ref_scene_set_time(hours)
        s.set_pose(pos, hor, ver)
        s.set_time(hours)
        s.update_camera(w, h)
        assert ref.camera().tobytes() == s.camera.tobytes()
        assert objects_equal(ref.objects(), s.objects)
        assert ref.lights().tobytes() == s.lights.tobytes()
        assert ref.ambient().tobytes() == s.ambient.tobytes()


def test_synthetic_inputs():
    n = 6
    t = rt.terrain_objects(n)
    assert len(t) == 2 * n * n + 3 and t["type"][0] == PLANE and (t["type"][1:-2] == TRIANGLE).all()
    assert list(np.nonzero(t["light"])[0]) == [len(t) - 2, len(t) - 1]
    # every terrain triangle faces up under the reference's culling (kernel.cu:104): normal.y > 0
    e1 = t["size"][1:-2] - t["pos"][1:-2]
    e2 = t["third"][1:-2] - t["pos"][1:-2]
    assert (np.cross(e1, e2)[:, 1] > 0).all()
    assert (t["pos"][1:-2][:, 1] >= -10.001).all() and (t["pos"][1:-2][:, 1] <= 300.001).all()
    sky = rt.synthetic_sky(64, 32)
    assert sky.shape == (4, 32, 64, 4) and (sky[..., 3] == 255).all()
    assert (rt.synthetic_sky(64, 32) == sky).all() and len({sky[k].tobytes() for k in range(4)}) == 4
