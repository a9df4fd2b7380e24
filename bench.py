#!/usr/bin/env python3
"""Benchmark of the ray-tracing hot path (BASELINE.json metric: Mrays/s and frames/s).

Workload at any N (BASELINE.json configs[1], the 10 Grays/s target config): the Frozen Island view
(key 5) at 1920x1080 with FXAA on, at all four times of day -- one STEP = those four frames, on the
reference's own sky maps (oracle/_ref/backgrounds/*.png = reference backgrounds/, scene.cpp:626-632) when
they travelled with the snapshot, else on the procedural stand-in of the same size.
Frames are independent, so N GPUs run frame-parallel: every rank renders the same batch
(weak scaling, no data-path collective); `value` = rays of all ranks / max-over-ranks device time.

  value     device-resident: scenes/LBVHs/sky already in HBM, CUDA events around the 4 x (trace + FXAA)
            kernels of each step; L2 flushed (256 MiB write) between steps, outside the events.
  e2e       through the reference-facing boundary launchKernel() with HOST Object/Light arrays each frame
            and every finished frame copied back to page-locked host memory by the library's asynchronous
            read-back (rt_readback_async / rt_wait_slot, three frames in flight), wall clock around K steps.
  roofline  dominant kernel = the tracer.  It is FP32/issue bound, not HBM bound, so the
            contract's `roofline` object (HBM) is accompanied by `roofline_fp32` (counted flops vs measured
            FFMA peak), `roofline_issue` and `roofline_fxaa` (the HBM-bound post-pass).
  sustained the same steps back to back for >= 1.5 s with its own clock record (the K timed steps are a burst).
  ref_kernel_sm100   the reference's OWN two kernels (kernel.cu:455-460) rebuilt for sm_100
            (oracle/_ref/libref_kernel_sm100.so) on the same four frames on this GPU: the column the north_star asks
            for next to ours.  A baseline leg like cpu_baseline: outside every timed region of the product.
  strong    ONE 3840x2160 frame of config 3 (Mountains, raised sea) and of config 4 (1M-triangle terrain) split
            over the N ranks by interleaved screen tiles, every rank's FXAA kernel storing its tiles straight into
            rank 0's frame over NVLink (peer frame): ms per composited frame, per-rank kernel split, and whether
            the composite equals the single-GPU frame.  This is the curve that moves with N; `value` is weak scaling.
  d2h_copy_gbs   a bare pinned device->host copy of one step's frames at this N: the ceiling `e2e` runs into.
  cpu_baseline / --impl reference: the reference's own kernel.cu text compiled for the host
            (oracle/_ref/libref_trace_cpu.so, kind "reference"; else the oracle port) on all host cores: the four
            full frames, FXAA included, scene states from the reference's own scene.cpp (oracle/_ref/libref_scene.so).

Ray count (the unit of work) is the reference algorithm's: every trace<d<=4> call plus every shadow ray
built at reference kernel.cu:182, counted once per frame by the instrumented kernel outside the timed region.
"""
import argparse
import ctypes as C
import hashlib
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

W, H = 1920, 1080
VIEW = 5
TODS = ("morning", "afternoon", "evening", "night")
TOD_HOURS = {"morning": 6.0, "afternoon": 14.0, "evening": 18.0, "night": 1.0}     # reference scene.cpp:713-728
POSE5 = ((-56.0, 2.2, 72.0), 309.0, -7.07)                                         # reference scene.cpp:736-741
SKY_W, SKY_H = 8192, 4096
SKY_DIR = os.path.join(ROOT, "oracle", "_ref", "backgrounds")
WORKLOAD = "configs[1]: Frozen Island view (key 5) 1920x1080 FXAA on, 4 times of day per step"
STRONG_TILE = (256, 128)          # screen tiles of the single-frame split: multiples of the 64x32 FXAA tile

# flop model of SURVEY 8d per test, by exit (reference kernel.cu:100-123, 49-66, 73-87)
F_BOX, F_PLANE, F_SHADE, F_SKY = 20, 21, 140, 40
F_TRI_CULLED, F_TRI_U, F_TRI_V, F_TRI_FULL, F_TRI_ACCEPT = 20, 30, 45, 51, 25
F_SPH_MISS, F_SPH_HIT = 16, 38


def config_dict(sky_label, world):
    """Identical in both arms: what is rendered, and how the GPU arm is timed."""
    return {"workload": WORKLOAD, "view": "key 5 (Frozen Island)", "size": "%dx%d" % (W, H), "fxaa": True,
            "hours": [TOD_HOURS[t] for t in TODS], "frames_per_step": 4, "sea_level": -4.5,
            "sky": ("reference backgrounds/*.png" if sky_label == "reference sky maps" else "procedural stand-in") +
                   " %dx%d x4" % (SKY_W, SKY_H),
            "parallelism": "frame-parallel x%d (GPU arm)" % world,
            "l2": "GPU arm: flushed by a 256 MiB write between steps"}


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.isfile(p):
        d = json.load(open(p))
        return d.get("hbm_gbs", 6650.0), "measured"
    return 6650.0, "fallback"


def csrc_hash():
    """sha256 over the kernel sources: profiles/traffic.json (ncu counters) is only valid for the sources it was
    captured from."""
    h = hashlib.sha256()
    d = os.path.join(ROOT, "raytracing_cuda_b200", "csrc")
    for fn in sorted(os.listdir(d)):
        if fn.endswith((".cu", ".cuh", ".h", ".cpp")) or fn == "Makefile":
            h.update(fn.encode())
            h.update(open(os.path.join(d, fn), "rb").read())
    return h.hexdigest()[:16]


class ClockSampler:
    """SM clocks / throttle reasons DURING a timed region (B200_PROFILING.md): NVML in-process every ~2 ms,
    nvidia-smi as the fallback."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index, uuid=None):
        self.gpu = gpu_index
        self.uuid = uuid
        self.proc = None
        self.lines = []
        self.nvml = None
        self.samples = []          # (sm_mhz, power_w, reasons bit mask) from NVML
        self.stop_flag = False

    def _nvml_loop(self):
        n = self.nvml
        while not self.stop_flag:
            try:
                sm = n.nvmlDeviceGetClockInfo(self.handle, n.NVML_CLOCK_SM)
                try:
                    rs = n.nvmlDeviceGetCurrentClocksEventReasons(self.handle)
                except Exception:
                    rs = n.nvmlDeviceGetCurrentClocksThrottleReasons(self.handle)
                try:
                    pw = n.nvmlDeviceGetPowerUsage(self.handle) / 1000.0
                except Exception:
                    pw = 0.0
                self.samples.append((float(sm), pw, int(rs)))
            except Exception:
                pass
            time.sleep(0.002)

    def start(self):
        try:
            import pynvml as n
            n.nvmlInit()
            self.handle = None
            if self.uuid:
                for cand in ("GPU-" + str(self.uuid), str(self.uuid)):
                    try:
                        self.handle = n.nvmlDeviceGetHandleByUUID(cand.encode() if hasattr(cand, "encode") else cand)
                        break
                    except Exception:
                        self.handle = None
            if self.handle is None:
                self.handle = n.nvmlDeviceGetHandleByIndex(self.gpu)
            self.max_mhz = float(n.nvmlDeviceGetMaxClockInfo(self.handle, n.NVML_CLOCK_SM))
            self.nvml = n
            self.thread = threading.Thread(target=self._nvml_loop, daemon=True)
            self.thread.start()
            return
        except Exception:
            self.nvml = None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.gpu), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "100"], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if self.nvml is not None:
            self.stop_flag = True
            self.thread.join(timeout=1.0)
            n = self.nvml
            bits = {"hw_slowdown": getattr(n, "nvmlClocksEventReasonHwSlowdown", 0x8),
                    "hw_thermal_slowdown": getattr(n, "nvmlClocksEventReasonHwThermalSlowdown", 0x40),
                    "sw_thermal_slowdown": getattr(n, "nvmlClocksEventReasonSwThermalSlowdown", 0x20),
                    "sw_power_cap": getattr(n, "nvmlClocksEventReasonSwPowerCap", 0x4)}
            sm = [x[0] for x in self.samples]
            pw = [x[1] for x in self.samples]
            reasons = sorted(k for k, b in bits.items() if any(x[2] & b for x in self.samples))
            return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": self.max_mhz if sm else None,
                    "reasons": reasons, "samples": len(sm), "power_w_max": max(pw) if pw else None,
                    "source": "NVML, 2 ms period"}
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=3)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1]))
                mx.append(float(f[2]))
            except ValueError:
                continue
            for name, val in zip(names, f[5:9]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


# ------------------------------------------------------------------------------------ CPU legs
# Nothing below this line up to run_reference_arm() loads the product library: scene states come from the reference's own
# scene.cpp (oracle/_ref/libref_scene.so) when it was built, the sky maps are decoded with Pillow.

def cpu_use_all_cores():
    """The CPU legs run on all host threads: torchrun exports OMP_NUM_THREADS=1 to its workers, which would put the
    reference's trace loop on one core.  Returns the thread count in use."""
    n = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)
    os.environ["OMP_NUM_THREADS"] = str(n)       # for a runtime that has not started yet
    try:
        gomp = C.CDLL("libgomp.so.1")            # the runtime the oracle / reference builds link
        gomp.omp_set_num_threads(C.c_int(n))
        gomp.omp_get_max_threads.restype = C.c_int
        return int(gomp.omp_get_max_threads())
    except Exception:
        return n


def cpu_sky():
    """(maps, label) for the CPU legs without the product library: the reference's PNGs through Pillow."""
    names = ("morning", "day", "evening", "night")
    if all(os.path.isfile(os.path.join(SKY_DIR, n + ".png")) for n in names):
        from PIL import Image
        Image.MAX_IMAGE_PIXELS = None
        maps = [np.asarray(Image.open(os.path.join(SKY_DIR, n + ".png")).convert("RGBA"), dtype=np.uint8) for n in names]
        return np.ascontiguousarray(np.stack(maps)), "reference sky maps"
    return None, "synthetic"


def cpu_states_from_reference_scene():
    """[(tod, objects, lights, frame)] from the reference's OWN scene.cpp (globals behind libref_scene.so), or None."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import oracle_util as ou
    from raytracing_cuda_b200.structs import Frame            # plain ctypes layouts; loads no library
    if ou.ref_lib("scene") is None:
        return None
    ref = ou.RefScene()
    out = []
    for tod in TODS:
        pos, hor, ver = POSE5
        ref.lib.ref_scene_set_aspect(W, H)
        ref.lib.ref_scene_set_pose(*pos, hor, ver)
        ref.lib.ref_scene_set_sea(-4.5)
        ref.lib.ref_scene_set_time(TOD_HOURS[tod])
        f = Frame()
        cam = ref.camera()
        C.memmove(C.byref(f.cam), cam.tobytes(), C.sizeof(f.cam))
        for k in range(3):
            f.ambient[k] = float(ref.ambient()[k])
        for k in range(4):
            f.skyVars[k] = float(ref.sky_vars()[k])
        f.dayTime = float(np.float32(TOD_HOURS[tod]) / np.float32(24.0))      # dayProgress, reference scene.cpp:822
        f.alias, f.imgw, f.imgh = 1, W, H
        out.append((tod, ref.objects(), ref.lights(), f))
    return out


def cpu_reference_step(states, sky):
    """One step on the host cores: the four full frames, FXAA included.  Returns (seconds, kind)."""
    import oracle_util as ou
    kind = "reference" if ou.ref_lib("trace_cpu") is not None else "port"
    t0 = time.perf_counter()
    for _, objs, lights, fr in states:
        if kind == "reference":
            ou.ref_cpu_render(objs, lights, fr, sky)
        else:
            ou.oracle_render(objs, lights, fr, sky)
    return time.perf_counter() - t0, kind


def cpu_step_rays(states, sky):
    """Rays of one step as the reference algorithm casts them, counted by the oracle (CPU arithmetic)."""
    import oracle_util as ou
    total = 0
    for _, objs, lights, fr in states:
        c = ou.oracle_render(objs, lights, fr, sky, want_out=False)["counters"]
        total += c["rays_primary"] + c["rays_shadow"] + c["rays_mirror"]
    return total


def run_reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    world = int(os.environ.get("WORLD_SIZE", "1"))
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    cores = cpu_use_all_cores()
    states = cpu_states_from_reference_scene()
    scene_src = "reference scene.cpp (oracle/_ref/libref_scene.so)"
    sky, sky_label = cpu_sky()
    if states is None or sky is None:
        # the reference builds did not travel: fall back to the host scene layer / procedural maps of the product library
        import raytracing_cuda_b200 as rt
        from raytracing_cuda_b200 import configs
        if states is None:
            states = configs.config2()
            scene_src = "restated scene layer (libref_scene.so absent)"
        if sky is None:
            sky = rt.synthetic_sky(SKY_W, SKY_H)
    rays = cpu_step_rays(states, sky)            # also warms the page cache of the sky maps
    for _ in range(args.warmup):
        cpu_reference_step(states, sky)
    total, kind = 0.0, "port"
    for _ in range(args.steps):
        dt, kind = cpu_reference_step(states, sky)
        total += dt
    ms = 1e3 * total / max(args.steps, 1)
    value = rays / (ms * 1e-3) / 1e6
    sample = "the 4 full frames of a step, FXAA included (%d rays per step); scene states: %s" % (rays, scene_src)
    line = {"impl": "reference", "metric": "Mrays/s", "value": value, "unit": "Mrays/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32",
            "data": "reference scene + " + sky_label,
            "config": config_dict(sky_label, world), "rays_per_step": rays, "timing": "host wall clock",
            "frames_per_s": 4.0 / (ms * 1e-3),
            "cpu_baseline": {"value": value, "unit": "Mrays/s", "cores": cores, "kind": kind, "sample": sample},
            "e2e": {"value": value, "unit": "Mrays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    try:        # evidence that this arm ran without the product library
        line["product_library_loaded"] = "libraytracing_b200" in open("/proc/self/maps").read()
    except OSError:
        pass
    print(json.dumps(line), flush=True)
    return 0


# ------------------------------------------------------------------------------------ GPU arm

def flops_of(c):
    """Counted flops of a frame from the instrumented kernel's counters, every test charged by its exit."""
    tri_culled = c["tri_tests"] - c["tri_exit_u"] - c["tri_exit_v"] - c["tri_full"]
    shaded = c["rays_shadow"] / 2.0
    return (c["node_tests"] * F_BOX + tri_culled * F_TRI_CULLED + c["tri_exit_u"] * F_TRI_U + c["tri_exit_v"] * F_TRI_V +
            c["tri_full"] * F_TRI_FULL + c["tri_accept"] * F_TRI_ACCEPT +
            (c["sphere_tests"] - c["sphere_hits"]) * F_SPH_MISS + c["sphere_hits"] * F_SPH_HIT +
            c["plane_tests"] * F_PLANE + shaded * F_SHADE + c["sky_lookups"] * F_SKY)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-extras", action="store_true", help="skip sustained / ref_kernel_sm100 / strong / d2h legs")
    ap.add_argument("--sky", default="auto", help="'auto' (the reference's PNGs when present), 'synthetic', or a directory")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference_arm(args)
    # stdout carries the one JSON line only: whatever libraries write to file descriptor 1 meanwhile (NCCL prints
    # its version banner there at N >= 4) goes to stderr; the line itself is written to the saved descriptor
    sys.stdout.flush()
    json_fd = os.dup(1)
    os.dup2(2, 1)

    import torch
    import raytracing_cuda_b200 as rt
    from raytracing_cuda_b200 import _lib, configs, sharding

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the product path has no CPU fallback")
    torch.cuda.set_device(local)
    dist = None
    if world > 1:
        import torch.distributed as dist
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")   # stdout carries the one JSON line only
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x):
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        if dist is not None:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    lib = _lib.load()
    states = configs.config2(W, H)
    sky_dir = None if args.sky == "synthetic" else (SKY_DIR if args.sky == "auto" else args.sky)
    sky, sky_label = configs.sky_maps(sky_dir)
    stream = torch.cuda.current_stream()

    # one context per time of day: scene + LBVH resident in HBM, sky uploaded once and shared
    ctxs = []
    for i, (tod, objs, lights, fr) in enumerate(states):
        r = rt.Renderer(local)
        if i == 0:
            r.set_sky(sky)
        else:
            r.share_sky(ctxs[0][0])
        r.set_scene(objs, lights)
        ctxs.append((r, fr))
    out_dev = torch.empty(W * H, dtype=torch.int32, device="cuda")

    # ray / test counts per frame (instrumented kernel, outside the timed region)
    rays_step, flops_step, counts = 0, 0.0, {}
    for r, fr in ctxs:
        r.set_option(_lib.RT_OPT_COUNTERS, 1)
        r.render(fr, out_dev)
        c = r.counters()
        r.set_option(_lib.RT_OPT_COUNTERS, 0)
        rays_step += rt.rays_per_frame(c)
        flops_step += flops_of(c)
        for k, v in c.items():
            counts[k] = counts.get(k, 0) + v
    fp32_peak = ctxs[0][0].fp32_peak_tflops()

    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")

    def step_device():
        for r, fr in ctxs:
            r.render(fr, out_dev)

    for _ in range(max(args.warmup, 3)):
        flush.fill_(1)
        step_device()
    barrier()
    try:
        dev_uuid = str(torch.cuda.get_device_properties(local).uuid)
    except Exception:
        dev_uuid = None
    sampler = ClockSampler(local, dev_uuid)
    sampler.start()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    launches0 = lib.rt_launch_count()
    t_wall0 = time.perf_counter()
    for k in range(args.steps):
        flush.fill_(k & 255)                      # L2 flush between timed iterations, outside the events
        ev[k][0].record(stream)
        step_device()
        ev[k][1].record(stream)
    torch.cuda.synchronize()
    t_wall = time.perf_counter() - t_wall0
    gpu_launches = int(lib.rt_launch_count() - launches0)          # counted by the library's launchers
    # per-kernel split: the same K steps once more with the library's own CUDA events around the tracer
    # (binning pre-pass + trace kernel) and the FXAA kernel, on the launching stream; averaged over the steps
    trace_ms = fxaa_ms = 0.0
    for r, _ in ctxs:
        r.set_option(_lib.RT_OPT_TIMING, 1)
    for k in range(args.steps):
        flush.fill_(k & 255)
        for r, fr in ctxs:
            r.render(fr, out_dev)
            r.synchronize()
            t = r.timings()
            trace_ms += t["trace_ms"] / args.steps
            fxaa_ms += t["fxaa_ms"] / args.steps
    for r, _ in ctxs:
        r.set_option(_lib.RT_OPT_TIMING, 0)
    clocks = sampler.stop()
    dev_ms = sum(a.elapsed_time(b) for a, b in ev)
    dev_ms_max = max_over_ranks(dev_ms)
    barrier()
    ms_per_step = dev_ms_max / args.steps
    value = world * rays_step / (ms_per_step * 1e-3) / 1e6

    # ---- e2e through launchKernel with host inputs and the library's asynchronous read-back of every frame
    from raytracing_cuda_b200.structs import Camera, Float3, Light, Object
    depth = max(2, min(_lib.RT_MAX_SLOTS, int(os.environ.get("RTB_E2E_DEPTH", "3"))))   # frames in flight
    pinned = rt.PinnedFrames(depth, W * H)
    dev2 = [torch.empty(W * H, dtype=torch.int32, device="cuda") for _ in range(depth)]
    sky_ptrs = [sky[k].ctypes.data for k in range(4)]
    legacy = [None]

    def legacy_frame(i, objs, lights, fr):
        slot = i % depth
        cam = Camera()
        C.memmove(C.byref(cam), C.byref(fr.cam), C.sizeof(Camera))
        skyv = (C.c_float * 4)(*fr.skyVars)
        lib.launchKernel(dev2[slot].data_ptr(), W, H, cam, C.cast(objs.ctypes.data, C.POINTER(Object)),
                         C.cast(lights.ctypes.data, C.POINTER(Light)), Float3(*fr.ambient), skyv, sky_ptrs[0],
                         sky_ptrs[1], sky_ptrs[2], sky_ptrs[3], sky.shape[2], sky.shape[1], fr.dayTime, True)
        if legacy[0] is None:
            legacy[0] = rt.Renderer.legacy()
        legacy[0].readback_async(dev2[slot], W * H, pinned[slot], slot)   # queued behind the frame, no host sync

    in_flight = []

    def step_e2e(counter):
        # `depth` frames in flight: frame i is traced while earlier frames are read back, the host runs ahead by the
        # rest; the pipeline keeps flowing from one step into the next and is drained inside the timed region
        for _, objs, lights, fr in states:
            if len(in_flight) >= depth:
                legacy[0].wait_slot(in_flight.pop(0))     # the buffers about to be reused have been read back
            legacy_frame(counter[0], objs, lights, fr)
            in_flight.append(counter[0] % depth)
            counter[0] += 1

    def drain():
        while in_flight:
            legacy[0].wait_slot(in_flight.pop(0))

    cnt = [0]
    for _ in range(max(args.warmup, 3)):
        step_e2e(cnt)
    drain()
    barrier()
    launches0 = lib.rt_launch_count()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        step_e2e(cnt)
    drain()
    torch.cuda.synchronize()
    e2e_s = max_over_ranks(time.perf_counter() - t0)
    e2e_launches = int(lib.rt_launch_count() - launches0)
    barrier()
    e2e_value = world * rays_step * args.steps / e2e_s / 1e6
    h2d = 4 * (133 * 68 + 2 * 28 + 16 + 12 + 4 + 108)
    d2h = 4 * W * H * 4

    extras = {}
    if not args.no_extras:
        # ---- bare pinned device -> host copy of one step's frames at this N (all ranks at once): e2e's ceiling
        cs = torch.cuda.Stream()
        host_t = [torch.empty(W * H, dtype=torch.int32).pin_memory() for _ in range(4)]
        barrier()
        for rep in range(2):                      # first pass warms up
            barrier()
            t0 = time.perf_counter()
            with torch.cuda.stream(cs):
                for _ in range(8):
                    for k in range(4):
                        host_t[k].copy_(dev2[k % depth], non_blocking=True)
            cs.synchronize()
            copy_s = max_over_ranks(time.perf_counter() - t0)
        extras["d2h_copy_gbs"] = {"value": world * 8 * d2h / copy_s / 1e9, "per_gpu": 8 * d2h / copy_s / 1e9,
                                  "what": "bare cudaMemcpyAsync of %d x 4 frames (%d B each) to pinned memory, all %d ranks at "
                                          "once, wall clock, max over ranks" % (8, 4 * W * H, world),
                                  "e2e_copy_gbs": world * d2h * args.steps / e2e_s / 1e9}

        # ---- sustained: the same steps back to back for >= 1.5 s, own clock record
        n_sus = max(args.steps, int(1500.0 / max(ms_per_step, 1e-3)) + 1)
        sus_sampler = ClockSampler(local, dev_uuid)
        barrier()
        sus_sampler.start()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.perf_counter()
        a.record(stream)
        for _ in range(n_sus):
            step_device()
        b.record(stream)
        torch.cuda.synchronize()
        sus_wall = time.perf_counter() - t0
        sus_clocks = sus_sampler.stop()
        sus_ms = max_over_ranks(a.elapsed_time(b))
        extras["sustained"] = {"value": world * rays_step * n_sus / (sus_ms * 1e-3) / 1e6, "unit": "Mrays/s",
                               "steps": n_sus, "seconds": sus_ms * 1e-3, "wall_seconds": sus_wall,
                               "ms_per_step": sus_ms / n_sus, "clocks": sus_clocks,
                               "what": "steps back to back on the stream, one event pair around all of them, no L2 flush "
                                       "(4 x 8192x4096 sky maps = 512 MiB exceed the 126 MB L2; frames stay in it)"}

        # ---- the reference's own kernels rebuilt for sm_100 on the same four frames (baseline leg, rank 0 only)
        if rank == 0:
            extras["ref_kernel_sm100"] = ref_kernel_leg(torch, states, sky, rays_step, ms_per_step)
        barrier()

        # ---- strong scaling: ONE 4K frame of configs 3 and 4 over the N ranks
        extras["strong"] = strong_leg(torch, rt, _lib, configs, sharding, ctxs[0][0], local, rank, world, dist, barrier,
                                      max_over_ranks)

    # ---- rooflines
    hbm_peak, peak_kind = peaks()
    npx = W * H
    sky_bytes = counts.get("sky_lookups", 0) * 4
    trace_bytes = 4 * npx * 4 + sky_bytes                      # framebuffer write + sky texels per step
    trace_s = max(trace_ms, 1e-6) * 1e-3
    fxaa_s = max(fxaa_ms, 1e-6) * 1e-3
    # ncu counters per launch (dram bytes, warp instructions) from the committed --set full captures: valid only for
    # the kernel sources they were captured from
    traffic, traffic_note = {}, None
    tp = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.isfile(tp):
        traffic = json.load(open(tp))
        if traffic.get("csrc_sha16") != csrc_hash():
            traffic_note = "profiles/traffic.json was captured from other kernel sources (%s, built %s): not used" % (
                traffic.get("csrc_sha16"), csrc_hash())
            traffic = {}
    else:
        traffic_note = "profiles/traffic.json absent"
    kname = "rt_trace_small_kernel (+ rt_cull_tiles_kernel)"
    roofline = {"bound": "hbm", "kernel": kname, "achieved": trace_bytes / trace_s / 1e9,
                "peak": hbm_peak, "unit": "GB/s", "frac": trace_bytes / trace_s / 1e9 / hbm_peak,
                "traffic": traffic.get("trace_bytes_per_step"), "traffic_note": traffic_note,
                "peak_source": peak_kind + " (MEASURED_PEAKS.json hbm_gbs)",
                "note": "tracer is FP32/issue bound (SURVEY 8d): see roofline_fp32 / roofline_issue; HBM-bound kernel is FXAA"}
    roofline_fp32 = {"bound": "fp32", "kernel": kname, "achieved": flops_step / trace_s / 1e12,
                     "peak": fp32_peak, "unit": "TFLOP/s", "frac": flops_step / trace_s / 1e12 / max(fp32_peak, 1e-9),
                     "flop_model": "per test by exit: box 20; triangle 20 culled / 30 u / 45 v / 51 full (+25 hit); sphere 16 / "
                                   "38; plane 21; shade 140; sky 40 (SURVEY 8d)",
                     "peak_source": "measured FFMA microbenchmark (rt_measure_fp32_peak); nominal 74.4"}
    winst = traffic.get("warp_inst_per_launch", {})
    roofline_issue = None
    if winst and clocks.get("sm_mhz"):
        per_step = 4.0 * (winst.get("rt_trace_small_kernel", 0) + winst.get("rt_cull_tiles_kernel", 0))
        n_sm = torch.cuda.get_device_properties(local).multi_processor_count
        issue_peak = n_sm * 4 * clocks["sm_mhz"] * 1e6
        roofline_issue = {"bound": "issue", "kernel": kname,
                          "achieved": per_step / trace_s / 1e9, "peak": issue_peak / 1e9, "unit": "Gwarp-inst/s",
                          "frac": per_step / trace_s / issue_peak,
                          "source": "warp instructions: ncu smsp__inst_executed.sum (profiles/traffic.json, sources %s); peak: "
                                    "SMs x 4 schedulers x SM clock under load" % traffic.get("csrc_sha16")}
    elif traffic_note:
        roofline_issue = {"bound": "issue", "frac": None, "note": traffic_note}
    roofline_fxaa = {"bound": "hbm", "kernel": "rt_fxaa_tma_kernel", "achieved": 8.0 * npx * 4 / fxaa_s / 1e9,
                     "peak": hbm_peak, "unit": "GB/s", "frac": 8.0 * npx * 4 / fxaa_s / 1e9 / hbm_peak,
                     "traffic": traffic.get("fxaa_bytes_per_step")}

    line = {"metric": "Mrays/s", "value": value, "unit": "Mrays/s", "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "reference scene + " + sky_label,
            "config": config_dict(sky_label, world), "rays_per_step": rays_step,
            "rays_per_pixel": rays_step / (4.0 * npx),
            "frames_per_s": world * 4 / (ms_per_step * 1e-3), "wall_ms_per_step_incl_flush": 1e3 * t_wall / args.steps,
            "kernel_ms_per_step": {"trace": trace_ms, "fxaa": fxaa_ms},
            "roofline": roofline, "roofline_fp32": roofline_fp32, "roofline_issue": roofline_issue,
            "roofline_fxaa": roofline_fxaa,
            "e2e": {"value": e2e_value, "unit": "Mrays/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                    "api": "launchKernel (host Object/Light arrays per frame) + rt_readback_async to page-locked host "
                           "memory, %d frames in flight" % depth,
                    "frames_per_s": world * 4 * args.steps / e2e_s, "gpu_launches": world * e2e_launches},
            "gpu_launches": world * gpu_launches, "gpu_launches_source": "rt_launch_count(): counted by the library's launchers",
            "clocks": clocks, "counters_per_step": counts}
    line.update(extras)

    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        sys.path.insert(0, os.path.join(ROOT, "tests"))
        cpu_cores = cpu_use_all_cores()
        cstates = cpu_states_from_reference_scene() or states
        rays = cpu_step_rays(cstates, sky)
        dt, reps, kind = 0.0, 0, "port"
        while dt < 10.0 and reps < 200:           # about 10 s of CPU work on the box's host cores
            d, kind = cpu_reference_step(cstates, sky)
            dt += d
            reps += 1
        line["cpu_baseline"] = {"value": reps * rays / dt / 1e6, "unit": "Mrays/s", "cores": cpu_cores, "kind": kind,
                                "sample": "%d x the 4 full frames of a step, FXAA included (%d rays each), %.1f s" % (
                                    reps, rays, dt)}
    if rank == 0:
        sys.stdout.flush()
        os.write(json_fd, (json.dumps(line) + "\n").encode())
    if dist is not None:
        dist.destroy_process_group()
    return 0


def ref_kernel_leg(torch, states, sky, rays_step, our_ms_per_step):
    """The reference's two kernels (kernel.cu:455-460), rebuilt for sm_100 by oracle/build_ref.py, timed on the same four
    frames with CUDA events (its own ref_time_kernels: `iters` back-to-back frames).  Baseline only."""
    path = os.path.join(ROOT, "oracle", "_ref", "libref_kernel_sm100.so")
    if not os.path.isfile(path):
        return {"available": False, "why": "oracle/_ref/libref_kernel_sm100.so did not travel (needs /root/reference at build time)"}
    from raytracing_cuda_b200 import _lib
    from raytracing_cuda_b200.structs import Camera, Float3, Light, Object
    ref = C.CDLL(path, mode=C.RTLD_LOCAL)
    ref.launchKernel.argtypes = _lib.SIGNATURES["launchKernel"][1]
    ref.launchKernel.restype = None
    ref.ref_time_kernels.argtypes = [C.c_void_p, C.c_int, C.c_int, Camera, C.c_bool, C.c_int]
    ref.ref_time_kernels.restype = C.c_float
    out = torch.empty(W * H, dtype=torch.int32, device="cuda")
    per = {}
    total_ms = 0.0
    for tod, objs, lights, fr in states:
        cam = Camera()
        C.memmove(C.byref(cam), C.byref(fr.cam), C.sizeof(Camera))
        ref.launchKernel(out.data_ptr(), W, H, cam, C.cast(objs.ctypes.data, C.POINTER(Object)),
                         C.cast(lights.ctypes.data, C.POINTER(Light)), Float3(*fr.ambient), (C.c_float * 4)(*fr.skyVars),
                         sky[0].ctypes.data, sky[1].ctypes.data, sky[2].ctypes.data, sky[3].ctypes.data, sky.shape[2],
                         sky.shape[1], fr.dayTime, True)          # uploads this frame's constants (and the sky, once)
        torch.cuda.synchronize()
        ref.ref_time_kernels(out.data_ptr(), W, H, cam, True, 3)
        ms = float(ref.ref_time_kernels(out.data_ptr(), W, H, cam, True, 20))
        per[tod] = ms
        total_ms += ms
    return {"available": True, "ms_per_frame": per, "ms_per_step": total_ms, "value": rays_step / (total_ms * 1e-3) / 1e6,
            "unit": "Mrays/s", "ours_over_reference_kernel": total_ms / our_ms_per_step,
            "what": "reference raytracing + antialiasing kernels (kernel.cu:455-460) rebuilt for sm_100, --use_fast_math, "
                    "20 back-to-back frames per time of day, warm L2"}


def strong_leg(torch, rt, _lib, configs, sharding, sky_owner, local, rank, world, dist, barrier, max_over_ranks):
    """ONE 3840x2160 frame of config 3 (Mountains, sea level 3.5) and of config 4 (terrain) over the N ranks:
    interleaved 256x128 screen tiles + 1-pixel apron, every rank's FXAA kernel stores its tiles into rank 0's frame
    over NVLink (rt_render_peer: hand-shake inside the kernel), rank 0 waits for all ranks and releases the buffer."""
    w, h = 3840, 2160
    tw, th = STRONG_TILE
    out = {"tile": "%dx%d" % (tw, th), "compose": "peer stores into rank 0's frame over NVLink (rt_render_peer), no collective"}
    r = rt.Renderer(local)
    r.share_sky(sky_owner)
    full = torch.empty(w * h, dtype=torch.int32, device="cuda")
    pf = None
    seq = [0]
    reps = 20

    def timed(fn):
        for _ in range(3):
            fn()
        barrier()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(reps):
            fn()
        b.record()
        torch.cuda.synchronize()
        ms = max_over_ranks(a.elapsed_time(b) / reps)
        barrier()
        return ms

    for name in ("config3", "config4"):
        if name == "config3":
            _, objs, lights, f = configs.config3(sea_levels=(3.5,))[0]
            what = "Mountains 3840x2160 FXAA on, sea level 3.5"
        else:
            objs, lights, f = configs.config4()
            what = "terrain %d objects 3840x2160 FXAA on" % len(objs)
            r.set_option(_lib.RT_OPT_SHADOW_BINS, 2)      # static scene: build the light grid with the first frame
        r.set_scene(objs, lights)
        r.set_option(_lib.RT_OPT_COUNTERS, 1)
        r.render(f, full)
        rays = rt.rays_per_frame(r.counters())
        r.set_option(_lib.RT_OPT_COUNTERS, 0)
        r.render(f, full)                                 # the single-GPU frame every rank holds for the comparison
        r.synchronize()
        res = {"what": what, "rays_per_frame": rays}
        res["ms_per_frame_one_gpu"] = timed(lambda: r.render(f, full))
        if world == 1:
            res["ms_per_frame"] = res["ms_per_frame_one_gpu"]
            res["composite_equals_single_gpu_frame"] = True
        else:
            if pf is None:
                def exchange(hb):
                    box = [hb]
                    dist.broadcast_object_list(box, src=0)
                    return box[0]
                pf = sharding.PeerFrame(r, rank, world, w * h, exchange)
            shard = rt.make_shard(world, rank, tw, th)

            def frame():
                seq[0] += 1
                pf.render(f, shard, seq[0])
                if rank == 0:
                    pf.wait(seq[0], check=False)
                    pf.release(seq[0])
            res["ms_per_frame"] = timed(frame)
            # one more composited frame into a zeroed buffer, compared with the single-GPU frame on rank 0
            seq[0] += 1
            if rank == 0:
                pf.tensor(seq[0]).zero_()
            barrier()
            pf.render(f, shard, seq[0])
            ok = 1.0
            if rank == 0:
                pf.wait(seq[0], check=False)
                r.synchronize()
                ok = 1.0 if (pf.error() == 0 and bool(torch.equal(pf.tensor(seq[0]), full))) else 0.0
                pf.release(seq[0])
            barrier()
            res["composite_equals_single_gpu_frame"] = bool(max_over_ranks(1.0 - ok) == 0.0)
            # this rank's kernel split for its shard (library events, one frame with a synchronisation)
            r.set_option(_lib.RT_OPT_TIMING, 1)
            seq[0] += 1
            pf.render(f, shard, seq[0])
            if rank == 0:
                pf.wait(seq[0], check=False)
                pf.release(seq[0])
            r.synchronize()
            t = r.timings()
            r.set_option(_lib.RT_OPT_TIMING, 0)
            res["shard_us_max_over_ranks"] = {"prepass_plus_trace": 1e3 * max_over_ranks(t["trace_ms"]),
                                              "fxaa_incl_acquire_and_peer_stores": 1e3 * max_over_ranks(t["fxaa_ms"])}
            barrier()
        res["frames_per_s"] = 1e3 / res["ms_per_frame"]
        res["Mrays_per_s"] = rays / res["ms_per_frame"] / 1e3
        res["speedup_vs_one_gpu"] = res["ms_per_frame_one_gpu"] / res["ms_per_frame"]
        out[name] = res
        if name == "config4":
            r.set_option(_lib.RT_OPT_SHADOW_BINS, 0)
    if pf is not None:
        barrier()
        pf.close()
    r.close()
    return out


if __name__ == "__main__":
    sys.exit(main())
