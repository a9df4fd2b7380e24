#!/usr/bin/env python3
"""Parity report on a B200: this library vs (a) the reference's own kernel.cu rebuilt for sm_100
(oracle/_ref/libref_kernel_sm100.so) and (b) the CPU oracle, for both views x 4 times of day.
Writes a table (markdown) to stdout / the given file.  TEST INFRASTRUCTURE (uses oracle/)."""
import ctypes as C
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import torch  # noqa: E402

import oracle_util as ou  # noqa: E402
import raytracing_cuda_b200 as rt  # noqa: E402
from raytracing_cuda_b200 import _lib  # noqa: E402
from raytracing_cuda_b200.structs import Camera, Float3, Light, Object  # noqa: E402


def ref_gpu_lib():
    lib = ou.ref_lib("kernel_sm100")
    if lib is None:
        return None
    lib.launchKernel.argtypes = _lib.SIGNATURES["launchKernel"][1]
    lib.launchKernel.restype = None
    lib.ref_set_hit_buffer.argtypes = [C.c_void_p, C.c_int]
    lib.ref_fxaa.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int]
    lib.ref_time_kernels.argtypes = [C.c_void_p, C.c_int, C.c_int, Camera, C.c_bool, C.c_int]
    lib.ref_time_kernels.restype = C.c_float
    return lib


def legacy_call(lib, out_ptr, s, f, sky):
    objs, lights = s.objects, s.lights
    cam = Camera()
    C.memmove(C.byref(cam), C.byref(f.cam), C.sizeof(Camera))
    skyv = (C.c_float * 4)(*f.skyVars)
    lib.launchKernel(out_ptr, f.imgw, f.imgh, cam, C.cast(objs.ctypes.data, C.POINTER(Object)),
                     C.cast(lights.ctypes.data, C.POINTER(Light)), Float3(*f.ambient), skyv, sky[0].ctypes.data,
                     sky[1].ctypes.data, sky[2].ctypes.data, sky[3].ctypes.data, sky.shape[2], sky.shape[1], f.dayTime,
                     bool(f.alias))
    return cam


def stats(a, b):
    m = ou.compare_images(a, b)
    ra, rb = ou.unpack_rgb(a), ou.unpack_rgb(b)
    m["n_gt1"] = int((np.abs(ra - rb).max(-1) > 1).sum())
    return m


def main():
    w, h = (int(sys.argv[1]), int(sys.argv[2])) if len(sys.argv) > 2 else (1280, 720)
    out_path = sys.argv[3] if len(sys.argv) > 3 else None
    do_cpu = w * h <= 1280 * 720
    ref = ref_gpu_lib()
    sky = rt.synthetic_sky(2048, 1024)
    r = rt.Renderer(0)
    r.set_sky(sky)
    s = rt.Scene()
    out = torch.zeros(w * h, dtype=torch.int32, device="cuda")
    hit = torch.full((w * h,), -2, dtype=torch.int32, device="cuda")
    if ref is not None:
        ref.ref_set_hit_buffer(hit.data_ptr(), w)
    lines = ["| view | time | AA | vs | ids differ (px) | exact px | within 1 LSB | >1 LSB (px) | max abs | PSNR dB |",
             "|---|---|---|---|---|---|---|---|---|---|"]
    for view in (5, 6):
        for tod, hours in rt.TIME_PRESETS.items():
            for alias in (False, True):
                f = s.configure(view=view, hours=hours, imgw=w, imgh=h, alias=alias)
                r.set_scene(s.objects, s.lights)
                img, ids = r.render_to_host(f, want_ids=True)
                if ref is not None:
                    legacy_call(ref, out.data_ptr(), s, f, sky)
                    torch.cuda.synchronize()
                    rimg = out.cpu().numpy().view(np.uint32).reshape(h, w).copy()
                    rids = hit.cpu().numpy().reshape(h, w).copy()
                    m = stats(img, rimg)
                    lines.append("| %d | %s | %s | reference kernel (sm_100) | %d | %.5f | %.5f | %d | %d | %.2f |" % (
                        view, tod, "on" if alias else "off", int((ids != rids).sum()), m["frac_exact"],
                        m["frac_within_1lsb"], m["n_gt1"], m["max_abs"], m["psnr_db"]))
                if do_cpu:
                    o = ou.oracle_render(s.objects, s.lights, f, sky)
                    m = stats(img, o["out"])
                    lines.append("| %d | %s | %s | CPU oracle | %d | %.5f | %.5f | %d | %d | %.2f |" % (
                        view, tod, "on" if alias else "off", int((ids != o["ids"]).sum()), m["frac_exact"],
                        m["frac_within_1lsb"], m["n_gt1"], m["max_abs"], m["psnr_db"]))
    # FXAA alone on identical input: ours (TMA and smem) vs the reference kernel
    if ref is not None:
        rng = np.random.default_rng(1)
        noise = rng.integers(0, 1 << 24, size=(h, w), dtype=np.uint32)
        f = s.configure(view=5, hours=18.0, imgw=w, imgh=h, alias=False)
        r.set_scene(s.objects, s.lights)
        frame = r.render_to_host(f)
        for name, src in (("noise", noise), ("frame v5 evening", frame)):
            pad = w + 4
            big = torch.zeros(w * h + 2 * pad, dtype=torch.int32, device="cuda")
            big[pad:pad + w * h] = torch.from_numpy(src.view(np.int32).reshape(-1)).cuda()
            src_dev = big[pad:pad + w * h]
            dst = torch.zeros(w * h, dtype=torch.int32, device="cuda")
            ref.ref_fxaa(src_dev.data_ptr(), dst.data_ptr(), w, h)
            torch.cuda.synchronize()
            want = dst.cpu().numpy().view(np.uint32).reshape(h, w).copy()
            for impl in (0, 1):
                r.set_option(_lib.RT_OPT_FXAA_IMPL, impl)
                src2 = src_dev.clone()
                r.fxaa(src2, dst, w, h)
                r.synchronize()
                got = dst.cpu().numpy().view(np.uint32).reshape(h, w)
                m = stats(got, want)
                lines.append("| - | FXAA only: %s | %s | reference kernel (sm_100) | - | %.6f | %.6f | %d | %d | %.2f |" % (
                    name, "tma" if impl == 0 else "smem", m["frac_exact"], m["frac_within_1lsb"], m["n_gt1"],
                    m["max_abs"], m["psnr_db"]))
            r.set_option(_lib.RT_OPT_FXAA_IMPL, 0)
    text = "Parity at %dx%d (procedural 2048x1024 sky)\n\n" % (w, h) + "\n".join(lines) + "\n"
    print(text)
    if out_path:
        with open(out_path, "w") as fh:
            fh.write(text)


if __name__ == "__main__":
    main()
