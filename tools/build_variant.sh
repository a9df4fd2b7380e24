#!/bin/bash
# Build an experimental variant of the library next to the product one:
#   tools/build_variant.sh <name> <extra nvcc flags...>   ->  build/variants/lib_<name>.so   (use with RTB_LIB_PATH)
set -e
cd "$(dirname "$0")/.."
name=$1; shift
out=build/variants; mkdir -p $out/obj_$name
NV="/usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -Xcompiler -fPIC"
for f in trace trace_small trace_big fxaa light_grid; do $NV --use_fast_math "$@" -c raytracing_cuda_b200/csrc/$f.cu -o $out/obj_$name/$f.o & done
for f in bvh_build api peak peer; do $NV "$@" -c raytracing_cuda_b200/csrc/$f.cu -o $out/obj_$name/$f.o & done
wait
$NV -shared -o $out/lib_$name.so $out/obj_$name/*.o build/obj/scene_host.o build/obj/synth.o build/obj/image_out.o build/obj/ingest.o -lz -Xcompiler -fopenmp -lgomp
echo $out/lib_$name.so
