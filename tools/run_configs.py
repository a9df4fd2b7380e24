#!/usr/bin/env python3
"""The five BASELINE.json configs, one JSON line each.  Single process or under torchrun
(one rank per GPU):

    python tools/run_configs.py [--configs 1,2,3,4,5] [--out profiles/r01_configs_n1.jsonl]
    python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 tools/run_configs.py ...

  1  Mountains view (key 6), 14:00, 1280x720, AA off, one frame (+ parity vs the CPU oracle when it is built)
  2  Island view (key 5), 1920x1080, FXAA on, the four times of day (= bench.py's workload)
  3  Mountains 3840x2160, FXAA on, sea-level sweep, ONE frame tile-sharded over the ranks + gather to rank 0
  4  synthetic terrain (708x708 cells = 1,002,528 triangles + plane + sun/moon), 3840x2160, tile-sharded;
     LBVH build timed separately
  5  256-frame camera path (key-5 pose -> key-6 pose, 24 h of automatic time), 1920x1080, frame i -> rank i % N

Times are device times (CUDA events on the launching stream), max over ranks; L2 is not flushed here
(bench.py does that for the headline number)."""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

import raytracing_cuda_b200 as rt  # noqa: E402
from raytracing_cuda_b200 import _lib, sharding  # noqa: E402

# screen tiles of the single-frame shards: large at 4K so that the 1-pixel apron stays cheap (13 % more patches at
# 256x128 against 56 % at 64x32), whole multiples of the 64x32 FXAA tile so that the TMA FXAA kernel applies
TILE_W, TILE_H = (int(v) for v in os.environ.get("RTB_TILE", "256x128").split("x"))
rank = int(os.environ.get("RANK", "0"))
world = int(os.environ.get("WORLD_SIZE", "1"))
local = int(os.environ.get("LOCAL_RANK", "0"))
dist = None


def barrier():
    if dist is not None:
        dist.barrier()
    torch.cuda.synchronize()


def max_over_ranks(x):
    t = torch.tensor([x], dtype=torch.float64, device="cuda")
    if dist is not None:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def sum_over_ranks(x):
    t = torch.tensor([x], dtype=torch.float64, device="cuda")
    if dist is not None:
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
    return float(t.item())


def timed(fn, reps, warm=3):
    for _ in range(warm):
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


def count_rays(r, f, out):
    r.set_option(_lib.RT_OPT_COUNTERS, 1)
    r.render(f, out)
    c = r.counters()
    r.set_option(_lib.RT_OPT_COUNTERS, 0)
    return rt.rays_per_frame(c), c


def emit(fh, d):
    d.update({"n_gpus": world, "gpu": torch.cuda.get_device_name(0)})
    if rank == 0:
        line = json.dumps(d)
        print(line, flush=True)
        if fh:
            fh.write(line + "\n")
            fh.flush()


def oracle_available():
    try:
        sys.path.insert(0, os.path.join(ROOT, "tests"))
        import oracle_util as ou
        return ou if os.path.isfile(os.path.join(ROOT, "oracle", "liboracle.so")) else None
    except Exception:
        return None


# ---------------------------------------------------------------------------------------------------

def config1(r, sky, fh):
    w, h = 1280, 720
    s = rt.Scene()
    f = s.configure(view=6, hours=14.0, imgw=w, imgh=h, alias=False)
    r.set_scene(s.objects, s.lights)
    out = torch.empty(w * h, dtype=torch.int32, device="cuda")
    rays, _ = count_rays(r, f, out)
    ms = timed(lambda: r.render(f, out), 100)
    d = {"config": 1, "what": "Mountains (key 6) 14:00 1280x720 AA off, one frame (every rank renders it)",
         "ms_per_frame": ms, "frames_per_s": 1e3 / ms, "Mrays_per_s": rays / ms / 1e3, "rays_per_frame": rays}
    ou = oracle_available()
    if ou is not None and rank == 0:
        img, ids = r.render_to_host(f, want_ids=True)
        t0 = time.perf_counter()
        o = ou.oracle_render(s.objects, s.lights, f, sky)
        d["cpu_oracle_s"] = time.perf_counter() - t0
        d["cpu_cores"] = os.cpu_count()
        m = ou.compare_images(img, o["out"])
        d["parity_vs_cpu_oracle"] = {"hit_id_mismatch_px": int((ids != o["ids"]).sum()),
                                     "frac_within_1lsb": m["frac_within_1lsb"], "psnr_db": m["psnr_db"]}
    emit(fh, d)


def config2(r, sky, fh):
    w, h = 1920, 1080
    s = rt.Scene()
    out = torch.empty(w * h, dtype=torch.int32, device="cuda")
    per = {}
    for tod in ("morning", "afternoon", "evening", "night"):
        f = s.configure(view=5, hours=rt.TIME_PRESETS[tod], imgw=w, imgh=h, alias=True)
        r.set_scene(s.objects, s.lights)
        rays, _ = count_rays(r, f, out)
        ms = timed(lambda: r.render(f, out), 100)
        per[tod] = {"ms_per_frame": ms, "Mrays_per_s": rays / ms / 1e3, "rays_per_frame": rays}
    tot_ms = sum(v["ms_per_frame"] for v in per.values())
    tot_rays = sum(v["rays_per_frame"] for v in per.values())
    emit(fh, {"config": 2, "what": "Island (key 5) 1920x1080 FXAA on, 4 times of day, frame-parallel (every rank "
              "renders all four)", "per_time_of_day": per, "Mrays_per_s": world * tot_rays / tot_ms / 1e3,
              "frames_per_s": world * 4e3 / tot_ms})


COMPOSE = os.environ.get("RTB_COMPOSE", "peer")     # peer: stores into rank 0's frame over NVLink; gather: NCCL
_peer = {"pf": None, "seq": 0}


def peer_frame(r, n_pixels):
    if _peer["pf"] is None:
        def exchange(hb):
            box = [hb]
            dist.broadcast_object_list(box, src=0)
            return box[0]
        _peer["pf"] = sharding.PeerFrame(r, rank, world, n_pixels, exchange)
    return _peer["pf"]


def sharded_frame(r, f, w, h, out_local, packed, frame0, bufs, shard):
    """Trace + FXAA this rank's tiles and composite the frame on rank 0: either every rank's FXAA kernel stores
    into rank 0's peer-mapped frame (flag hand-shake, no collective), or pack + NCCL gather + unpack."""
    if world == 1:
        r.render(f, out_local, shard=shard)
        return
    if COMPOSE == "peer":
        pf = peer_frame(r, w * h)
        _peer["seq"] += 1
        seq = _peer["seq"]
        pf.render(f, shard, seq)
        if rank == 0:
            pf.wait(seq, check=False)
            _peer["last"] = seq
            pf.release(seq)          # (a consumer would use pf.tensor(seq) between wait and release)
        return
    r.render(f, out_local, shard=shard)
    r.pack_tiles(out_local, w, h, shard, packed)
    dist.gather(packed, gather_list=bufs if rank == 0 else None, dst=0)
    if rank == 0:
        for k in range(world):
            r.unpack_tiles(bufs[k], w, h, rt.make_shard(world, k, shard.tile_w, shard.tile_h), frame0)


def composite_ok(r, f, w, h, out_local, packed, frame0, bufs, shard, full):
    """One more composited frame, compared on rank 0 with the unsharded frame `full`."""
    if world == 1:
        return True
    if COMPOSE == "peer":
        pf = peer_frame(r, w * h)
        _peer["seq"] += 1
        seq = _peer["seq"]
        pf.tensor(seq).zero_() if rank == 0 else None
        barrier()
        pf.render(f, shard, seq)
        ok = True
        if rank == 0:
            pf.wait(seq, check=False)
            torch.cuda.synchronize()
            ok = bool(torch.equal(pf.tensor(seq), full)) and pf.error() == 0
            pf.release(seq)
        barrier()
        return ok
    frame0.zero_()
    sharded_frame(r, f, w, h, out_local, packed, frame0, bufs, shard)
    torch.cuda.synchronize()
    return bool((frame0 == full).all().item()) if rank == 0 else True


def config3(r, sky, fh):
    w, h = 3840, 2160
    s = rt.Scene()
    shard = rt.make_shard(world, rank, TILE_W, TILE_H) if world > 1 else None
    out_local = torch.empty(w * h, dtype=torch.int32, device="cuda")
    frame0 = torch.zeros(w * h, dtype=torch.int32, device="cuda")
    packed = bufs = None
    if world > 1:
        n_slots = r.tile_slots(shard, w, h)
        packed = torch.zeros(n_slots * TILE_W * TILE_H, dtype=torch.int32, device="cuda")
        bufs = [torch.zeros_like(packed) for _ in range(world)] if rank == 0 else None
    res = []
    for sea in (-4.5, -2.5, -0.5, 1.5, 3.5, 5.5, 7.5, 9.5, 11.5):
        s.set_sea_level(sea)
        f = s.configure(view=6, hours=14.0, imgw=w, imgh=h, alias=True)
        r.set_scene(s.objects, s.lights)
        rays, _ = count_rays(r, f, out_local)              # whole frame, unsharded
        full = out_local.clone()
        ms = timed(lambda: sharded_frame(r, f, w, h, out_local, packed, frame0, bufs, shard), 20)
        ok = composite_ok(r, f, w, h, out_local, packed, frame0, bufs, shard, full)
        res.append({"sea_level": sea, "ms_per_frame": ms, "Mrays_per_s": rays / ms / 1e3, "rays_per_frame": rays,
                    "composite_equals_single_gpu_frame": ok})
    ms_all = sum(x["ms_per_frame"] for x in res)
    emit(fh, {"config": 3, "what": "Mountains 3840x2160 FXAA on, sea-level sweep; one frame tile-sharded (%dx%d tiles, " % (TILE_W, TILE_H) +
              "1-px apron) over the ranks; composite on rank 0 by " + ("peer stores over NVLink (no collective)" if COMPOSE == "peer" else "pack + NCCL gather + unpack"),
              "sweep": res, "frames_per_s": len(res) * 1e3 / ms_all,
              "Mrays_per_s": sum(x["rays_per_frame"] for x in res) / ms_all / 1e3})


def config4(r, sky, fh):
    w, h = 3840, 2160
    cells = 708
    objs = rt.terrain_objects(cells)
    s = rt.Scene()
    s.set_pose((324.4, 330.0, -84.0), 141.2, 35.0)       # positive verAngle looks DOWN (reference scene.cpp:114)
    s.set_time(14.0)
    s.set_antialiasing(True)
    f = s.frame(w, h)
    lights = s.lights
    r.set_option(_lib.RT_OPT_TIMING, 1)
    r.set_scene(objs, lights)
    r.synchronize()
    build_ms = r.timings()["bvh_ms"]
    # a second full build (different object count forces one) for a warm number
    r.set_scene(objs[:-1], lights)
    r.set_scene(objs, lights)
    r.synchronize()
    build_ms_warm = r.timings()["bvh_ms"]
    r.set_option(_lib.RT_OPT_TIMING, 0)
    shard = rt.make_shard(world, rank, TILE_W, TILE_H) if world > 1 else None
    out_local = torch.empty(w * h, dtype=torch.int32, device="cuda")
    frame0 = torch.zeros(w * h, dtype=torch.int32, device="cuda")
    packed = bufs = None
    if world > 1:
        n_slots = r.tile_slots(shard, w, h)
        packed = torch.zeros(n_slots * TILE_W * TILE_H, dtype=torch.int32, device="cuda")
        bufs = [torch.zeros_like(packed) for _ in range(world)] if rank == 0 else None
    # the light grid of a big scene is built by the second frame that finds scene and lights unchanged: first frame
    # (traversal), then the frame that builds it (timed), then the counters with the grid in place
    r.render(f, out_local)
    r.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    r.render(f, out_local)
    e1.record()
    r.synchronize()
    grid_frame_ms = e0.elapsed_time(e1)
    rays, c = count_rays(r, f, out_local)
    full = out_local.clone()
    # the whole 4K frame once more with the shadow rays traversing the hierarchy: must be the same pixels
    r.set_option(_lib.RT_OPT_SHADOW_BINS, 1)
    r.render(f, out_local)
    r.synchronize()
    grid_ok = bool(torch.equal(out_local, full))
    t_trav = timed(lambda: r.render(f, out_local), 10)
    r.set_option(_lib.RT_OPT_SHADOW_BINS, 0)
    r.render(f, out_local)
    r.synchronize()
    ms = timed(lambda: sharded_frame(r, f, w, h, out_local, packed, frame0, bufs, shard), 20)
    ok = composite_ok(r, f, w, h, out_local, packed, frame0, bufs, shard, full)
    # cross-check against the brute-force kernel (reference arithmetic over all primitives) on a small frame
    fs = s.frame(96, 54)
    fs.alias = 0
    a, ia = r.render_to_host(fs, want_ids=True)
    r.set_option(_lib.RT_OPT_TRAVERSAL, 1)
    b, ib = r.render_to_host(fs, want_ids=True)
    r.set_option(_lib.RT_OPT_TRAVERSAL, 0)
    emit(fh, {"config": 4, "what": "synthetic terrain %dx%d cells = %d objects, 3840x2160 FXAA on, tile-sharded, composite on rank 0 (%s)"
              % (cells, cells, len(objs), COMPOSE), "ms_per_frame": ms, "frames_per_s": 1e3 / ms, "Mrays_per_s": rays / ms / 1e3,
              "rays_per_frame": rays, "lbvh_build_ms_first": build_ms, "lbvh_build_ms": build_ms_warm,
              "light_grid_build_ms_incl_one_whole_frame": grid_frame_ms,
              "light_grid_equals_traversal_whole_frame": grid_ok, "ms_per_whole_frame_traversal_only": t_trav,
              "node_tests_per_pixel": c["node_tests"] / (w * h), "tri_tests_per_pixel": c["tri_tests"] / (w * h),
              "composite_equals_single_gpu_frame": ok,
              "bvh_equals_bruteforce_96x54": bool((ia == ib).all() and (a == b).all())})


def config5(r, sky, fh):
    w, h = 1920, 1080
    n_frames = 256
    s = rt.Scene()
    (p5, h5, v5), (p6, h6, v6) = rt.POSES[5], rt.POSES[6]
    out = torch.empty(w * h, dtype=torch.int32, device="cuda")
    mine = sharding.frames_of_rank(n_frames, rank, world)
    states = []
    for i in mine:
        a = i / 255.0
        pos = tuple(p5[k] + a * (p6[k] - p5[k]) for k in range(3))
        s.set_pose(pos, h5 + a * (h6 - h5), v5 + a * (v6 - v5))
        s.set_time(float(np.fmod(np.float32(6.0) + np.float32(0.5 * 0.1875 * i), np.float32(24.0))))
        s.set_antialiasing(True)
        f = s.frame(w, h)
        fr = type(f)()
        import ctypes as C
        C.memmove(C.byref(fr), C.byref(f), C.sizeof(f))
        states.append((s.objects.copy(), s.lights.copy(), fr))
    rays = 0
    for objs, lights, fr in states:                      # ray count, outside the timing
        r.set_scene(objs, lights)
        rays += count_rays(r, fr, out)[0]
    rays = sum_over_ranks(rays)

    def batch():
        for objs, lights, fr in states:
            r.set_scene(objs, lights)                    # per-frame upload + LBVH refit (sun/moon, colours move)
            r.render(fr, out)
    ms = timed(batch, 3, warm=1)
    emit(fh, {"config": 5, "what": "256-frame camera path key-5 -> key-6 pose, automatic time of day (24 h), 1920x1080 "
              "FXAA on, frame i -> rank i % N; each frame = host scene upload + LBVH refit + trace + FXAA",
              "ms_per_batch": ms, "frames_per_s": n_frames * 1e3 / ms, "Mrays_per_s": rays / ms / 1e3,
              "rays_per_batch": rays})


def main():
    global dist
    ap = argparse.ArgumentParser()
    ap.add_argument("--configs", default="1,2,3,4,5")
    ap.add_argument("--out", default="")
    args = ap.parse_args()
    torch.cuda.set_device(local)
    if world > 1:
        import torch.distributed as d
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        d.init_process_group("nccl", device_id=torch.device("cuda", local))
        dist = d
    fh = open(args.out, "a") if (args.out and rank == 0) else None
    sky = rt.synthetic_sky(8192, 4096)
    r = rt.Renderer(local)
    r.set_sky(sky)
    fns = {1: config1, 2: config2, 3: config3, 4: config4, 5: config5}
    for c in [int(x) for x in args.configs.split(",")]:
        fns[c](r, sky, fh)
        barrier()
    if dist is not None:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
