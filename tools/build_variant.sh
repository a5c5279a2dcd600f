#!/bin/bash
# Build a tuning variant of librt_b200.so into build/variants/<name>/ (git-ignored; travels to
# the GPU box).  usage: tools/build_variant.sh <name> [extra nvcc flags...]
# trace_fast.cu and rt_accel_build.cu are recompiled with the flags (they are knobs of the fast policy's
# kernels, their launch code and the tree layout both sides must agree on); the other objects come from
# python_ray_tracer_b200/csrc (run make there first) -- so the exact policy of a variant that changes the
# tree layout does not match its builder: measure the fast policy only with those.
# Select it at run time with RT_B200_LIBRARY=build/variants/<name>/librt_b200.so
set -euo pipefail
name=$1; shift
root=$(cd "$(dirname "$0")/.." && pwd)
src=$root/python_ray_tracer_b200/csrc
out=$root/build/variants/$name
mkdir -p "$out"
NVCC=${NVCC:-/usr/local/cuda/bin/nvcc}
FLAGS="-gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo -Xcompiler -fPIC,-Wall,-Wno-unused-function --expt-relaxed-constexpr"
$NVCC $FLAGS "$@" -Xptxas -v -c "$src/trace_fast.cu" -o "$out/trace_fast.o" 2>&1 | grep -A2 "render_rows_small_kernelILb0ELi3ELi8\|render_rows_kernelILi1ELb0ELi128ELi1ELi8" | grep -E "Used|spill" | sed "s/^/[$name] /"
$NVCC $FLAGS "$@" -c "$src/rt_accel_build.cu" -o "$out/rt_accel_build.o"
$NVCC -gencode arch=compute_100a,code=sm_100a -shared -o "$out/librt_b200.so" "$out/trace_fast.o" "$out/rt_accel_build.o" \
  "$src"/{trace_exact,rt_abi,rt_abi_host,rt_peaks,rt_png_host,rt_png_dev}.o
