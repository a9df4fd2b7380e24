#!/usr/bin/env python3
"""profiles/traffic.json from an `ncu -i capture.ncu-rep --page raw --csv` export of tools/profile_one.py (Island 1920x1080):
per-launch DRAM bytes and warp instructions of the frame's three kernels (median over the captured launches), stamped
with the hash of the kernel sources the capture was taken from (bench.py uses the file only when the hash matches the
sources it runs).  usage: make_traffic.py raw.csv capture-name [out.json]"""
import csv
import json
import os
import statistics
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402

rows = list(csv.reader(open(sys.argv[1], errors="replace")))
hdr = rows[0]
units = rows[1]
col = {n: i for i, n in enumerate(hdr)}


def num(r, name):
    v = float(r[col[name]].replace(",", "") or 0)
    u = units[col[name]].lower()
    scale = {"byte": 1, "kbyte": 1e3, "mbyte": 1e6, "gbyte": 1e9, "ns": 1, "us": 1e3, "usecond": 1e3, "msecond": 1e6,
             "nsecond": 1, "ms": 1e6, "inst": 1, "": 1}.get(u, 1)
    return v * scale


kernels = {"rt_cull_tiles_kernel": [], "rt_trace_small_kernel": [], "rt_fxaa_tma_kernel": []}
for r in rows[2:]:
    if len(r) < len(hdr):
        continue
    name = r[col["Kernel Name"]]
    for k in kernels:
        if name.startswith(k) or k in name:
            kernels[k].append(r)
out = {"_comment": "dram__bytes_read.sum + dram__bytes_write.sum and smsp__inst_executed.sum per launch (median over the launches "
                   "of the capture) from the ncu --set full capture %s (Island 1920x1080 14:00, tools/profile_one.py; ncu starts "
                   "each kernel with cold caches), x4 frames per step." % sys.argv[2],
       "csrc_sha16": bench.csrc_hash(), "per_launch": {}, "warp_inst_per_launch": {"_source": "smsp__inst_executed.sum of the same capture"},
       "duration_ns_per_launch": {}}
for k, rs in kernels.items():
    if not rs:
        continue
    rd = statistics.median(num(r, "dram__bytes_read.sum") for r in rs)
    wr = statistics.median(num(r, "dram__bytes_write.sum") for r in rs)
    out["per_launch"][k] = {"read": int(rd), "write": int(wr), "launches": len(rs)}
    out["warp_inst_per_launch"][k] = int(statistics.median(num(r, "smsp__inst_executed.sum") for r in rs))
    out["duration_ns_per_launch"][k] = int(statistics.median(num(r, "gpu__time_duration.sum") for r in rs))
pl = out["per_launch"]
out["trace_bytes_per_step"] = 4 * sum(pl[k]["read"] + pl[k]["write"] for k in ("rt_cull_tiles_kernel", "rt_trace_small_kernel") if k in pl)
out["fxaa_bytes_per_step"] = 4 * sum(pl[k]["read"] + pl[k]["write"] for k in ("rt_fxaa_tma_kernel",) if k in pl)
path = sys.argv[3] if len(sys.argv) > 3 else os.path.join(ROOT, "profiles", "traffic.json")
json.dump(out, open(path, "w"), indent=1)
print(json.dumps(out, indent=1))
