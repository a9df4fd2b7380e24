#!/usr/bin/env python3
"""profiles/<name>_ncu_full.md from the files tools/ncu_capture.sh leaves in gpurun_out/: the key lines of
`ncu --page details` per kernel and the top source lines by warp instructions / stall samples (tools/ncu_by_line.py).
usage: make_ncu_summary.py TAG out.md"""
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
tag, out = sys.argv[1], sys.argv[2]
keep = re.compile(r"^\s+(Duration|Elapsed Cycles|SM Active Cycles|Memory Throughput|DRAM Throughput|L1/TEX Cache Throughput|"
                  r"L2 Cache Throughput|Compute \(SM\) Throughput|Executed Ipc Active|Executed Ipc Elapsed|Issue Slots Busy|"
                  r"Issued Ipc Active|Mem Busy|Max Bandwidth|L1/TEX Hit Rate|L2 Hit Rate|One or More Eligible|No Eligible|"
                  r"Registers Per Thread|Shared Memory Configuration Size|Dynamic Shared Memory Per Block|"
                  r"Static Shared Memory Per Block|Block Limit Registers|Block Limit Shared Mem|Theoretical Occupancy|"
                  r"Achieved Occupancy|Achieved Active Warps Per SM|Avg. Active Threads Per Warp|"
                  r"Avg. Not Predicated Off Threads Per Warp|Executed Instructions|Waves Per SM|Grid Size|Block Size)\s")
lines = ["# ncu --set full --clock-control none --import-source on: one Island 1920x1080 frame chain "
         "(tools/profile_one.py 5 14 1920 1080 8, launches 16-18), capture `%s`" % tag, "", "```"]
for ln in open(os.path.join(ROOT, "gpurun_out", tag + "_full_details.txt"), errors="replace"):
    if re.match(r"^  \S", ln) and "rt_" in ln and "(" in ln:
        lines.append(ln.rstrip()[:170])
    elif keep.match(ln):
        lines.append(ln.rstrip())
lines.append("```")
for kern in ("rt_trace_small_kernel", "rt_fxaa_tma_kernel", "rt_cull_tiles_kernel"):
    lines += ["", "## %s: source lines by warp instructions executed / stall samples" % kern, "", "```"]
    r = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "ncu_by_line.py"),
                        os.path.join(ROOT, "gpurun_out", tag + "_full_source.csv"), "22", kern],
                       capture_output=True, text=True)
    lines += [l[:190] for l in r.stdout.splitlines()]
    lines.append("```")
open(out, "w").write("\n".join(lines) + "\n")
print(out, len(lines), "lines")
