#!/usr/bin/env python3
"""Benchmark of the ray-tracing hot path (BASELINE.json metric: Mrays/s and frames/s).

Workload at any N (BASELINE.json configs[1], the 10 Grays/s target config): the Frozen Island view
(key 5) at 1920x1080 with FXAA on, at all four times of day -- one STEP = those four frames.
Frames are independent, so N GPUs run frame-parallel: every rank renders the same batch
(weak scaling, no data-path collective); `value` = rays of all ranks / max-over-ranks device time.

  value     device-resident: scenes/LBVHs/sky already in HBM, CUDA events around the 4 x (trace + FXAA)
            kernels of each step; L2 flushed (256 MiB write) between steps, outside the events.
  e2e       through the reference-facing boundary launchKernel() with HOST Object/Light arrays each frame
            and the finished frame copied back to pinned host memory, wall clock around K steps.
  roofline  dominant kernel = the tracer.  It is FP32/latency bound, not HBM bound, so the
            contract's `roofline` object (HBM) is accompanied by `roofline_fp32` (counted flops vs measured
            FFMA peak) and `roofline_fxaa` (the HBM-bound post-pass).
  cpu_baseline / --impl reference: the reference's own kernel.cu text compiled for the host
            (oracle/_ref/libref_trace_cpu.so, kind "reference"; else the oracle port) on all host cores,
            on a bounded sample (every 4th row of the same four frames).

Ray count (the unit of work) is the reference algorithm's: every trace<d<=4> call plus every shadow ray
built at reference kernel.cu:182, counted once per frame by the instrumented kernel outside the timed region.
"""
import argparse
import ctypes as C
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
SKY_W, SKY_H = 8192, 4096
WORKLOAD = "configs[1]: Frozen Island view (key 5) 1920x1080 FXAA on, 4 times of day per step"

# per-test flop constants of SURVEY 8d (reference kernel.cu:100-123, 49-66, 73-87)
F_BOX, F_TRI, F_SPH, F_PLANE, F_SHADE, F_SKY = 20, 45, 27, 21, 140, 40


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.isfile(p):
        d = json.load(open(p))
        return d.get("hbm_gbs", 6650.0), "measured"
    return 6650.0, "fallback"


class ClockSampler:
    """nvidia-smi clocks/throttle reasons DURING the timed region (B200_PROFILING.md)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index, uuid=None):
        self.gpu = gpu_index
        self.uuid = uuid
        self.proc = None
        self.lines = []
        self.nvml = None
        self.samples = []          # (sm_mhz, sm_max_mhz, reasons bit mask) from NVML
        self.stop_flag = False

    def _nvml_loop(self):
        # in-process NVML sampling every ~2 ms: the timed region is tens of milliseconds, an nvidia-smi process
        # (100 ms period, slower to start on an 8-GPU box) can miss it altogether
        n = self.nvml
        while not self.stop_flag:
            try:
                sm = n.nvmlDeviceGetClockInfo(self.handle, n.NVML_CLOCK_SM)
                try:
                    rs = n.nvmlDeviceGetCurrentClocksEventReasons(self.handle)
                except Exception:
                    rs = n.nvmlDeviceGetCurrentClocksThrottleReasons(self.handle)
                self.samples.append((float(sm), self.max_mhz, int(rs)))
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
            reasons = sorted(k for k, b in bits.items() if any(x[2] & b for x in self.samples))
            return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": self.max_mhz if sm else None,
                    "reasons": reasons, "samples": len(sm), "source": "NVML, 2 ms period"}
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


def scene_states():
    import raytracing_cuda_b200 as rt
    s = rt.Scene()
    states = []
    for tod in TODS:
        f = s.configure(view=VIEW, hours=rt.TIME_PRESETS[tod], imgw=W, imgh=H, alias=True)
        fr = type(f)()
        C.memmove(C.byref(fr), C.byref(f), C.sizeof(f))
        states.append((tod, s.objects.copy(), s.lights.copy(), fr))
    return states


# ------------------------------------------------------------------------------------ CPU legs

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


def cpu_reference_step(states, sky, y_step):
    """One bounded sample of the step on the host cores.  Returns (rays traced, seconds, kind)."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import oracle_util as ou
    kind = "reference" if ou.ref_lib("trace_cpu") is not None else "port"
    rays = 0
    t0 = time.perf_counter()
    for _, objs, lights, fr in states:
        if kind == "reference":
            ou.ref_cpu_render(objs, lights, fr, sky, y_step=y_step)
        else:
            ou.oracle_render(objs, lights, fr, sky, y_step=y_step, want_out=False)
    dt = time.perf_counter() - t0
    return dt, kind


def cpu_sample_rays(states, sky, y_step):
    """Ray count of the sampled rows, from the oracle's counters (same rows)."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import oracle_util as ou
    total = 0
    for _, objs, lights, fr in states:
        c = ou.oracle_render(objs, lights, fr, sky, y_step=y_step, want_out=False)["counters"]
        total += c["rays_primary"] + c["rays_shadow"] + c["rays_mirror"]
    return total


def run_reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    import raytracing_cuda_b200 as rt
    states = scene_states()
    sky = rt.synthetic_sky(SKY_W, SKY_H)
    y_step = 4
    cores = cpu_use_all_cores()
    rays = cpu_sample_rays(states, sky, y_step)          # also warms the page cache of the sky maps
    for _ in range(args.warmup):
        cpu_reference_step(states, sky, y_step)
    total = 0.0
    kind = "port"
    for _ in range(args.steps):
        dt, kind = cpu_reference_step(states, sky, y_step)
        total += dt
    ms = 1e3 * total / max(args.steps, 1)
    value = rays / (ms * 1e-3) / 1e6
    sample = "rows y%%%d==0 of the 4 frames (%d rays per step), FXAA skipped" % (y_step, rays)
    line = {"impl": "reference", "metric": "Mrays/s", "value": value, "unit": "Mrays/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": WORKLOAD, "sky": "procedural %dx%d x4" % (SKY_W, SKY_H), "timing": "host wall clock"},
            "cpu_baseline": {"value": value, "unit": "Mrays/s", "cores": cores, "kind": kind, "sample": sample},
            "e2e": {"value": value, "unit": "Mrays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)
    return 0


# ------------------------------------------------------------------------------------ GPU arm

def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--sky", default="synthetic", help="'synthetic' or a directory with the reference's backgrounds/*.png")
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
    from raytracing_cuda_b200 import _lib

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

    states = scene_states()
    sky = rt.synthetic_sky(SKY_W, SKY_H) if args.sky == "synthetic" else rt.load_sky_pngs(args.sky)
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
        shaded = c["rays_shadow"] / 2.0
        flops_step += (c["node_tests"] * F_BOX + c["tri_tests"] * F_TRI + c["sphere_tests"] * F_SPH +
                       c["plane_tests"] * F_PLANE + shaded * F_SHADE + c["sky_lookups"] * F_SKY)
        for k, v in c.items():
            counts[k] = counts.get(k, 0) + v
    fp32_peak = ctxs[0][0].fp32_peak_tflops()

    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")

    def step_device(timing):
        for r, fr in ctxs:
            r.set_option(_lib.RT_OPT_TIMING, 1 if timing else 0)
            r.render(fr, out_dev)

    for _ in range(max(args.warmup, 3)):
        flush.fill_(1)
        step_device(False)
    barrier()
    try:
        dev_uuid = str(torch.cuda.get_device_properties(local).uuid)
    except Exception:
        dev_uuid = None
    sampler = ClockSampler(local, dev_uuid)
    sampler.start()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    trace_ms = fxaa_ms = 0.0
    t_wall0 = time.perf_counter()
    for k in range(args.steps):
        flush.fill_(k & 255)                      # L2 flush between timed iterations, outside the events
        ev[k][0].record(stream)
        step_device(False)
        ev[k][1].record(stream)
    torch.cuda.synchronize()
    t_wall = time.perf_counter() - t_wall0
    # per-kernel split: the same K steps once more with the library's own CUDA events around the tracer
    # (binning pre-pass + trace kernel) and the FXAA kernel, on the launching stream; averaged over the steps
    for k in range(args.steps):
        flush.fill_(k & 255)
        for r, fr in ctxs:
            r.set_option(_lib.RT_OPT_TIMING, 1)
            r.render(fr, out_dev)
            r.synchronize()
            t = r.timings()
            trace_ms += t["trace_ms"] / args.steps
            fxaa_ms += t["fxaa_ms"] / args.steps
            r.set_option(_lib.RT_OPT_TIMING, 0)
    clocks = sampler.stop()
    dev_ms = sum(a.elapsed_time(b) for a, b in ev)
    t = torch.tensor([dev_ms], dtype=torch.float64, device="cuda")
    if dist is not None:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    barrier()
    dev_ms_max = float(t.item())
    ms_per_step = dev_ms_max / args.steps
    value = world * rays_step / (ms_per_step * 1e-3) / 1e6

    # ---- e2e through launchKernel with host inputs and a device->host read-back of every frame
    lib = _lib.load()
    from raytracing_cuda_b200.structs import Camera, Float3, Light, Object
    depth = max(2, int(os.environ.get("RTB_E2E_DEPTH", "3")))     # frames in flight (render + read-back pipeline)
    pinned = [torch.empty(W * H, dtype=torch.int32).pin_memory() for _ in range(depth)]
    dev2 = [torch.empty(W * H, dtype=torch.int32, device="cuda") for _ in range(depth)]
    copy_stream = torch.cuda.Stream()
    sky_ptrs = [sky[k].ctypes.data for k in range(4)]

    def legacy_frame(i, objs, lights, fr):
        cam = Camera()
        C.memmove(C.byref(cam), C.byref(fr.cam), C.sizeof(Camera))
        skyv = (C.c_float * 4)(*fr.skyVars)
        lib.launchKernel(dev2[i % depth].data_ptr(), W, H, cam, C.cast(objs.ctypes.data, C.POINTER(Object)),
                         C.cast(lights.ctypes.data, C.POINTER(Light)), Float3(*fr.ambient), skyv, sky_ptrs[0],
                         sky_ptrs[1], sky_ptrs[2], sky_ptrs[3], SKY_W if args.sky == "synthetic" else sky.shape[2],
                         SKY_H if args.sky == "synthetic" else sky.shape[1], fr.dayTime, True)
        done = torch.cuda.Event()
        done.record(torch.cuda.default_stream())
        copy_stream.wait_event(done)
        with torch.cuda.stream(copy_stream):
            pinned[i % depth].copy_(dev2[i % depth], non_blocking=True)
        copied = torch.cuda.Event()
        copied.record(copy_stream)
        return copied

    pending = []

    def step_e2e(counter):
        # `depth` frames in flight: frame i is traced while earlier frames are read back, the host runs ahead by the
        # rest; the pipeline keeps flowing from one step into the next and is drained inside the timed region
        for _, objs, lights, fr in states:
            if len(pending) >= depth:
                pending.pop(0).synchronize()      # the buffer about to be reused has been read back
            pending.append(legacy_frame(counter[0], objs, lights, fr))
            counter[0] += 1

    def drain():
        while pending:
            pending.pop(0).synchronize()

    cnt = [0]
    for _ in range(max(args.warmup, 3)):
        step_e2e(cnt)
    drain()
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        step_e2e(cnt)
    drain()
    torch.cuda.synchronize()
    e2e_s = time.perf_counter() - t0
    t = torch.tensor([e2e_s], dtype=torch.float64, device="cuda")
    if dist is not None:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    barrier()
    e2e_value = world * rays_step * args.steps / float(t.item()) / 1e6
    h2d = 4 * (133 * 68 + 2 * 28 + 16 + 12 + 4 + 108)
    d2h = 4 * W * H * 4

    # ---- rooflines (last step's kernel split)
    hbm_peak, peak_kind = peaks()
    npx = W * H
    sky_bytes = counts.get("sky_lookups", 0) * 4
    trace_bytes = 4 * npx * 4 + sky_bytes                      # framebuffer write + sky texels per step
    trace_s = max(trace_ms, 1e-6) * 1e-3
    fxaa_s = max(fxaa_ms, 1e-6) * 1e-3
    traffic = {}
    tp = os.path.join(ROOT, "profiles", "traffic.json")   # dram bytes per launch from the committed ncu --set full captures
    if os.path.isfile(tp):
        traffic = json.load(open(tp))
    roofline = {"bound": "hbm", "kernel": "rt_trace_small_kernel (+ rt_cull_tiles_kernel)", "achieved": trace_bytes / trace_s / 1e9,
                "peak": hbm_peak, "unit": "GB/s", "frac": trace_bytes / trace_s / 1e9 / hbm_peak,
                "traffic": traffic.get("trace_bytes_per_step"),
                "peak_source": peak_kind + " (MEASURED_PEAKS.json hbm_gbs)",
                "note": "tracer is FP32/latency bound (SURVEY 8d): see roofline_fp32; HBM-bound kernel is FXAA"}
    roofline_fp32 = {"bound": "fp32", "kernel": "rt_trace_small_kernel (+ rt_cull_tiles_kernel)", "achieved": flops_step / trace_s / 1e12,
                     "peak": fp32_peak, "unit": "TFLOP/s", "frac": flops_step / trace_s / 1e12 / max(fp32_peak, 1e-9),
                     "peak_source": "measured FFMA microbenchmark (rt_measure_fp32_peak); nominal 74.4"}
    # what actually bounds the tracer: warp-instruction issue (4 schedulers per SM, one instruction per clock each).
    # Instructions per launch from the committed ncu capture (static for this workload), time measured live.
    winst = traffic.get("warp_inst_per_launch", {})
    roofline_issue = None
    if winst and clocks.get("sm_mhz"):
        per_step = 4.0 * (winst.get("rt_trace_small_kernel", 0) + winst.get("rt_cull_tiles_kernel", 0))
        n_sm = torch.cuda.get_device_properties(local).multi_processor_count
        issue_peak = n_sm * 4 * clocks["sm_mhz"] * 1e6
        roofline_issue = {"bound": "issue", "kernel": "rt_trace_small_kernel (+ rt_cull_tiles_kernel)",
                          "achieved": per_step / trace_s / 1e9, "peak": issue_peak / 1e9, "unit": "Gwarp-inst/s",
                          "frac": per_step / trace_s / issue_peak,
                          "source": "warp instructions: ncu smsp__inst_executed.sum (profiles/traffic.json); peak: SMs x 4 schedulers x SM clock under load"}
    roofline_fxaa = {"bound": "hbm", "kernel": "rt_fxaa_tma_kernel", "achieved": 8.0 * npx * 4 / fxaa_s / 1e9,
                     "peak": hbm_peak, "unit": "GB/s", "frac": 8.0 * npx * 4 / fxaa_s / 1e9 / hbm_peak,
                     "traffic": traffic.get("fxaa_bytes_per_step")}

    line = {"metric": "Mrays/s", "value": value, "unit": "Mrays/s", "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic" if args.sky == "synthetic" else "reference sky maps",
            "config": {"workload": WORKLOAD, "frames_per_step": 4, "rays_per_step": rays_step,
                       "rays_per_pixel": rays_step / (4.0 * npx), "parallelism": "frame-parallel x%d" % world,
                       "sky": "procedural %dx%d x4" % (SKY_W, SKY_H) if args.sky == "synthetic" else args.sky,
                       "l2": "flushed by a 256 MiB write between steps"},
            "frames_per_s": world * 4 / (ms_per_step * 1e-3), "wall_ms_per_step_incl_flush": 1e3 * t_wall / args.steps,
            "kernel_ms_per_step": {"trace": trace_ms, "fxaa": fxaa_ms},
            "roofline": roofline, "roofline_fp32": roofline_fp32, "roofline_issue": roofline_issue,
            "roofline_fxaa": roofline_fxaa,
            "e2e": {"value": e2e_value, "unit": "Mrays/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                    "api": "launchKernel (host Object/Light arrays per frame) + frame copied to pinned host memory",
                    "frames_per_s": world * 4 * args.steps / float(t.item())},
            "gpu_launches": world * args.steps * 4 * 3,   # per frame: binning pre-pass, tracer, FXAA
            "clocks": clocks, "counters_per_step": counts}

    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        y_step = 4
        cpu_cores = cpu_use_all_cores()
        rays = cpu_sample_rays(states, sky, y_step)
        dt, reps, kind = 0.0, 0, "port"
        while dt < 10.0 and reps < 200:           # about 10 s of CPU work on the box's host cores
            d, kind = cpu_reference_step(states, sky, y_step)
            dt += d
            reps += 1
        line["cpu_baseline"] = {"value": reps * rays / dt / 1e6, "unit": "Mrays/s", "cores": cpu_cores,
                                "kind": kind,
                                "sample": "%d x rows y%%%d==0 of the 4 frames (%d rays each), FXAA skipped, %.1f s" % (
                                    reps, y_step, rays, dt)}
    if rank == 0:
        sys.stdout.flush()
        os.write(json_fd, (json.dumps(line) + "\n").encode())
    if dist is not None:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
