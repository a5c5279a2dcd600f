#!/bin/bash
# Build a tuning variant of librt_b200.so into build/variants/<name>/ (git-ignored; travels to
# the GPU box).  usage: tools/build_variant.sh <name> [extra nvcc flags...]
# Only trace_fast.cu is recompiled with the flags (they are knobs of the fast policy's kernels and
# their launch code); the other objects come from python_ray_tracer_b200/csrc (run make there first).
# Select it at run time with RT_B200_LIBRARY=build/variants/<name>/librt_b200.so
set -euo pipefail
name=$1; shift
root=$(cd "$(dirname "$0")/.." && pwd)
src=$root/python_ray_tracer_b200/csrc
out=$root/build/variants/$name
mkdir -p "$out"
NVCC=${NVCC:-/usr/local/cuda/bin/nvcc}
FLAGS="-gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo -Xcompiler -fPIC,-Wall,-Wno-unused-function --expt-relaxed-constexpr"
$NVCC $FLAGS "$@" -Xptxas -v -c "$src/trace_fast.cu" -o "$out/trace_fast.o" 2>&1 | grep -A2 "render_rows_small_kernelILb0ELi3ELi8" | grep -E "Used|spill" | sed "s/^/[$name] /"
$NVCC -gencode arch=compute_100a,code=sm_100a -shared -o "$out/librt_b200.so" "$out/trace_fast.o" \
  "$src"/{trace_exact,rt_abi,rt_abi_host,rt_peaks,rt_accel_build,rt_png_host,rt_png_dev}.o
