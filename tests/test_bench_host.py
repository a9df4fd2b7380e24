"""Host logic of bench.py that needs no GPU: the reference arm (the reference's own CPU code on the same config as
the GPU arm) and the bookkeeping the GPU arm's line is built from."""
import ctypes as C
import json
import os
import subprocess
import sys

import numpy as np
import pytest

import oracle_util as ou

sys.path.insert(0, ou.ROOT)
import bench  # noqa: E402
from raytracing_cuda_b200 import configs  # noqa: E402


@pytest.mark.skipif(ou.ref_lib("scene") is None, reason="oracle/_ref/libref_scene.so not built (needs /root/reference)")
def test_reference_arm_states_are_the_gpu_arms_states():
    """The reference arm takes Object / Light / Camera from the reference's own scene.cpp, the GPU arm from the
    restated scene layer: the two must describe the same four frames, field for field."""
    ref_states = bench.cpu_states_from_reference_scene()
    our_states = configs.config2(bench.W, bench.H)
    assert [s[0] for s in ref_states] == [s[0] for s in our_states] == list(bench.TODS)
    for (_, ro, rl, rf), (_, oo, ol, of) in zip(ref_states, our_states):
        for name in ("type", "shine", "specular", "mirror", "color", "pos", "size", "light"):
            assert (ro[name] == oo[name]).all(), name
        tri = ro["type"] == 2
        assert (ro["third"][tri] == oo["third"][tri]).all()      # `third` of spheres / planes is never initialised
        assert rl.tobytes() == ol.tobytes()
        assert bytes(rf.cam) == bytes(of.cam)
        assert list(rf.ambient) == list(of.ambient) and list(rf.skyVars) == list(of.skyVars)
        assert rf.dayTime == of.dayTime and (rf.alias, rf.imgw, rf.imgh) == (of.alias, of.imgw, of.imgh) == (1, 1920, 1080)


def test_config_dict_is_the_same_for_both_arms_and_hash_is_stable():
    assert bench.config_dict("reference sky maps", 2) == bench.config_dict("reference sky maps", 2)
    assert bench.config_dict("synthetic", 1)["sky"].startswith("procedural")
    assert len(bench.csrc_hash()) == 16 and bench.csrc_hash() == bench.csrc_hash()


def test_flop_model_charges_tests_by_exit():
    c = dict(node_tests=10, tri_tests=100, tri_exit_u=20, tri_exit_v=10, tri_full=5, tri_accept=3, sphere_tests=8,
             sphere_hits=2, plane_tests=4, rays_shadow=6, sky_lookups=1)
    want = 10 * 20 + 65 * 20 + 20 * 30 + 10 * 45 + 5 * 51 + 3 * 25 + 6 * 16 + 2 * 38 + 4 * 21 + 3 * 140 + 40
    assert bench.flops_of(c) == want


@pytest.mark.skipif(ou.ref_lib("trace_cpu") is None, reason="oracle/_ref/libref_trace_cpu.so not built")
def test_reference_arm_end_to_end_one_step():
    """`bench.py --impl reference`: one full step (4 frames at 1920x1080 with FXAA) of the reference's own trace text
    on the host cores; same config dict as the GPU arm; never loads the product library when the reference builds
    and sky maps are there."""
    r = subprocess.run([sys.executable, os.path.join(ou.ROOT, "bench.py"), "--impl", "reference", "--steps", "1",
                        "--warmup", "0"], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-2000:]
    line = json.loads(r.stdout.strip().splitlines()[-1])
    assert line["impl"] == "reference" and line["cpu_baseline"]["kind"] == "reference"
    have_sky = os.path.isdir(ou.REF_BACKGROUNDS)
    assert line["config"] == bench.config_dict("reference sky maps" if have_sky else "synthetic", 1)
    assert line["e2e"]["value"] == line["value"] > 0 and line["unit"] == "Mrays/s"
    assert 26e6 < line["rays_per_step"] < 28e6          # 3.27 rays per pixel x 4 frames
    if have_sky and ou.ref_lib("scene") is not None:
        assert line["product_library_loaded"] is False
