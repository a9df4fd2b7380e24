"""Parity with the ground truth the north_star names: the reference's OWN kernel.cu rebuilt for sm_100
(oracle/_ref/libref_kernel_sm100.so, oracle/build_ref.py), run on this GPU next to the product, on the BASELINE.json
configs at their own sizes and on the reference's own sky maps (oracle/_ref/backgrounds/*.png, scene.cpp:626-632;
the procedural stand-in of the same size when they did not travel).

Gates (BASELINE.json north_star): primary hit ids equal except near-tie pixels; 8-bit RGB within 1 LSB on >= 99.9 % of
pixels and PSNR >= 50 dB.  Everything the product renders here goes through its C-ABI.

The reference reads its textures on the FIRST launchKernel call only (kernel.cu:414-442) and sizes its scratch image
by the texture dimensions (kernel.cu:441): one module-wide sky of 8192x4096 texels serves every test below (the
largest frame, 3840x2160, has fewer pixels than that)."""
import ctypes as C
import os

import numpy as np
import pytest

import oracle_util as ou
import raytracing_cuda_b200 as rt
from raytracing_cuda_b200 import _lib, configs

pytestmark = pytest.mark.gpu

ID_MISMATCH_MAX = 2e-4
WITHIN_1LSB_MIN = 0.999
PSNR_MIN = 50.0


class _World:
    """Module-wide: the sky, the reference kernel library (initialised with it) and a product context."""

    def __init__(self):
        import torch
        self.torch = torch
        lib = ou.ref_lib("kernel_sm100")
        if lib is None:
            pytest.skip("oracle/_ref/libref_kernel_sm100.so did not travel")
        lib.launchKernel.argtypes = _lib.SIGNATURES["launchKernel"][1]
        lib.launchKernel.restype = None
        lib.ref_set_hit_buffer.argtypes = [C.c_void_p, C.c_int]
        self.ref = lib
        self.sky, self.sky_label = configs.sky_maps(ou.REF_BACKGROUNDS)
        self.r = rt.Renderer(0)
        self.r.set_sky(self.sky)
        self.stats = []

    def reference(self, objs, lights, f):
        """(image, primary hit ids) of the reference's two kernels for this state."""
        torch = self.torch
        w, h = f.imgw, f.imgh
        out = torch.zeros(w * h, dtype=torch.int32, device="cuda")
        hit = torch.full((w * h,), -2, dtype=torch.int32, device="cuda")
        self.ref.ref_set_hit_buffer(hit.data_ptr(), w)
        cam = rt.structs.Camera()
        C.memmove(C.byref(cam), C.byref(f.cam), C.sizeof(cam))
        objs = np.ascontiguousarray(objs)
        lights = np.ascontiguousarray(lights)
        sky = self.sky
        self.ref.launchKernel(out.data_ptr(), w, h, cam, C.cast(objs.ctypes.data, C.POINTER(rt.structs.Object)),
                              C.cast(lights.ctypes.data, C.POINTER(rt.structs.Light)), rt.structs.Float3(*f.ambient),
                              (C.c_float * 4)(*f.skyVars), sky[0].ctypes.data, sky[1].ctypes.data, sky[2].ctypes.data,
                              sky[3].ctypes.data, sky.shape[2], sky.shape[1], f.dayTime, bool(f.alias))
        torch.cuda.synchronize()
        self.ref.ref_set_hit_buffer(None, w)
        return (out.cpu().numpy().view(np.uint32).reshape(h, w).copy(), hit.cpu().numpy().reshape(h, w).copy())

    def product(self, objs, lights, f):
        self.r.set_scene(objs, lights)
        return self.r.render_to_host(f, want_ids=True)

    def gate(self, what, objs, lights, f):
        ref_img, ref_ids = self.reference(objs, lights, f)
        img, ids = self.product(objs, lights, f)
        m = ou.compare_images(img, ref_img)
        m["id_mismatch_px"] = int((ids != ref_ids).sum())
        self.stats.append("%-46s %dx%d %s" % (what, f.imgw, f.imgh, m))
        assert m["frac_within_1lsb"] >= WITHIN_1LSB_MIN, (what, m)
        assert m["psnr_db"] >= PSNR_MIN, (what, m)
        assert m["id_mismatch_px"] <= ID_MISMATCH_MAX * ids.size, (what, m)
        return m


@pytest.fixture(scope="module")
def world():
    w = _World()
    yield w
    w.r.close()
    os.makedirs(os.path.join(ou.ROOT, "gpurun_out"), exist_ok=True)
    with open(os.path.join(ou.ROOT, "gpurun_out", "parity_vs_reference_kernel.txt"), "w") as fh:
        fh.write("product vs the reference's kernel.cu rebuilt for sm_100, same GPU; sky: %s\n" % w.sky_label)
        for line in w.stats:
            fh.write(line + "\n")


def test_config1_mountains_720p_aa_off(world):
    objs, lights, f = configs.config1()
    world.gate("config 1 Mountains 14:00 AA off", objs, lights, f)


@pytest.mark.parametrize("view", [5, 6])
def test_both_views_720p_four_times_of_day(world, view):
    s = rt.Scene()
    for tod, hours in rt.TIME_PRESETS.items():
        f = s.configure(view=view, hours=hours, imgw=1280, imgh=720, alias=True)
        world.gate("view %d %s" % (view, tod), s.objects, s.lights, f)


def test_config2_island_1080p_four_times_of_day(world):
    """The benchmarked workload itself: every frame of a bench.py step."""
    for tod, objs, lights, f in configs.config2():
        assert (f.imgw, f.imgh, f.alias) == (1920, 1080, 1)
        world.gate("config 2 Island %s" % tod, objs, lights, f)


def test_config3_mountains_4k_sea_levels(world):
    for sea, objs, lights, f in configs.config3(sea_levels=(-4.5, 7.5)):
        assert (f.imgw, f.imgh) == (3840, 2160)
        world.gate("config 3 Mountains sea %+.1f" % sea, objs, lights, f)


def test_config5_path_frames_through_refit(world):
    """Nine of the 256 path frames -- the four plateaus, one frame inside each of the four cross-fade intervals
    (scene.cpp:788-803) and the last -- rendered IN ORDER on one context, so that every frame after the first takes
    rt_set_scene's refit path (sun / moon spheres moved, colours changed, scene.cpp:674-687, 758-776)."""
    frames = (0, 32, 64, 117, 140, 160, 200, 245, 255)
    hours = [configs.config5_time(i) for i in frames]
    assert 8 < hours[1] < 10 and 16 < hours[3] < 18 and 20 < hours[5] < 22 and 4 < hours[7] < 6
    for i, objs, lights, f in configs.config5(frames):
        world.gate("config 5 frame %3d (%.2f h)" % (i, configs.config5_time(i)), objs, lights, f)
    assert world.r.bvh_info()["prims"] == 132
