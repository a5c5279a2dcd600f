#!/bin/bash
# Round-2 variant sweep of the fast policy's kernels (tools/build_variant.sh builds, tools/sweep.sh runs).
#   here (no GPU):  tools/r02_sweep.sh build
#   on the box:     tools/r02_sweep.sh run > gpurun_out/r02_sweep.txt 2>&1
set -uo pipefail
root=$(cd "$(dirname "$0")/.." && pwd)
cd "$root"
case "${1:-}" in
build)
  rm -rf build/variants
  tools/build_variant.sh base
  tools/build_variant.sh mb6 -DRT_SMALL_MIN_BLOCKS=6                      # 80 registers, 6 blocks per SM
  tools/build_variant.sh mb4 -DRT_SMALL_MIN_BLOCKS=4                      # 128 registers, 4 blocks per SM
  tools/build_variant.sh single -DRT_CHUNK4_ROUNDS=0 -DRT_CHUNK2_ROUNDS=0  # one ticket per draw, as in round 1
  tools/build_variant.sh chunk2 -DRT_CHUNK4_ROUNDS=0 -DRT_CHUNK2_ROUNDS=2  # at most two tickets per draw
  tools/build_variant.sh chunk84 -DRT_CHUNK4_ROUNDS=8 -DRT_CHUNK2_ROUNDS=4 # chunks end earlier
  tools/build_variant.sh spec32 -DRT_SPEC_F32_A2=0.03f                    # rough materials: |H|^2 and N.H in FP32
  tools/build_variant.sh nosmall -DRT_SMALL_MAX=1                         # README scene through the staged-table kernel
  ;;
run)
  for wl in c2 c1 c4; do
    echo "== workload $wl"
    tools/sweep.sh --workload $wl --steps 100 --warmup 5
  done
  ;;
*) echo "usage: $0 build|run"; exit 2 ;;
esac
