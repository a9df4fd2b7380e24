#!/bin/bash
# On the GPU box: launch list of the bench command + ncu --set full capture of one Island 1080p frame chain, exported to
# CSV / text under gpurun_out/ (read with tools/ncu_by_line.py and tools/make_traffic.py).  usage: ncu_capture.sh TAG
TAG=${1:-r02}
set -x
python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-extras > gpurun_out/${TAG}_bench_short.json 2> gpurun_out/${TAG}_bench_short.err || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${TAG}_launches.csv \
    python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-extras > gpurun_out/${TAG}_ncu_bench.log 2>&1
python tools/profile_one.py 5 14 1920 1080 8 > gpurun_out/${TAG}_profile_one.log 2>&1 || exit 1
ncu --set full --clock-control none --import-source on -k regex:"rt_cull_tiles_kernel|rt_trace_small_kernel|rt_fxaa_tma_kernel" \
    --launch-skip 15 --launch-count 3 -f -o gpurun_out/${TAG}_full python tools/profile_one.py 5 14 1920 1080 8 > gpurun_out/${TAG}_ncu_full.log 2>&1
ncu -i gpurun_out/${TAG}_full.ncu-rep --page raw --csv > gpurun_out/${TAG}_full_raw.csv
ncu -i gpurun_out/${TAG}_full.ncu-rep --page details > gpurun_out/${TAG}_full_details.txt
ncu -i gpurun_out/${TAG}_full.ncu-rep --page source --csv --print-source sass,cuda > gpurun_out/${TAG}_full_source.csv 2>/dev/null
ls -la gpurun_out/${TAG}_*
