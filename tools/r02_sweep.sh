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
  make -C python_ray_tracer_b200/csrc >/dev/null
  (tools/build_variant.sh base) &
  (tools/build_variant.sh mb6 -DRT_SMALL_MIN_BLOCKS=6) &                       # 80 registers, 6 blocks per SM
  (tools/build_variant.sh single -DRT_CHUNK4_ROUNDS=0 -DRT_CHUNK2_ROUNDS=0) &  # one ticket per draw, as in round 1
  (tools/build_variant.sh chunk2 -DRT_CHUNK4_ROUNDS=0 -DRT_CHUNK2_ROUNDS=2) &  # pairs, single tickets for the last 2 rounds
  wait
  (tools/build_variant.sh chunk2b -DRT_CHUNK4_ROUNDS=0 -DRT_CHUNK2_ROUNDS=6) & # pairs end earlier
  (tools/build_variant.sh chunk42 -DRT_CHUNK4_ROUNDS=12 -DRT_CHUNK2_ROUNDS=4) & # fours only at the very start
  (tools/build_variant.sh mb6c2 -DRT_SMALL_MIN_BLOCKS=6 -DRT_CHUNK4_ROUNDS=0 -DRT_CHUNK2_ROUNDS=2) &
  (tools/build_variant.sh timing -DRT_TIMING_DEBUG -DRT_CHUNK4_ROUNDS=0 -DRT_CHUNK2_ROUNDS=2) &
  wait
  ;;
run)
  for wl in c2 c1 c4; do
    echo "== workload $wl"
    tools/sweep.sh --workload $wl --steps 100 --warmup 5
  done
  ;;
*) echo "usage: $0 build|run"; exit 2 ;;
esac
