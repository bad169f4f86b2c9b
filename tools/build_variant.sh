#!/bin/bash
# Tuning builds of libnm03.so with other compile-time shapes: tools/build_variant.sh <name> -DNM03_K1_TW=64 ...
# -> gpurun_variants/<name>.so (git-ignored, shipped to the GPU box); time with NM03_LIB=<that .so> python tools/k_time.py
set -e
name=$1; shift
root=$(cd "$(dirname "$0")/.." && pwd)
pkg=$root/nm03-capstone-project_b200
out=$root/gpurun_variants/$name.build
mkdir -p $out
flags="-gencode arch=compute_100a,code=sm_100a -std=c++17 -O3 -lineinfo -fmad=false -Xcompiler -fPIC $*"
pids=""
for w in 1 3 5 7 9 11 13 15; do
  nvcc $flags -DNM03_K1_WINDOW=$w -Xptxas -v -c $pkg/csrc/nm03_k1_w.cu -o $out/k1_w$w.o 2> $out/k1_w$w.log & pids="$pids $!"
done
nvcc $flags -c $pkg/csrc/nm03_kernels.cu -o $out/kernels.o & pids="$pids $!"
nvcc $flags -c $pkg/csrc/nm03_api.cu -o $out/api.o & pids="$pids $!"
for p in $pids; do wait $p; done
nvcc -gencode arch=compute_100a,code=sm_100a -shared -o $root/gpurun_variants/$name.so $out/*.o -lpthread
grep -h "Used" $out/k1_w7.log | sort | uniq -c | head -3
rm -rf $out
