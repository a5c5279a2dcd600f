#!/bin/bash
# First GPU call of the next round: the experiments staged at the end of round 1,
# each a build variant (tools/build_variant.sh), parity first, then device times.
#   here (no GPU):  tools/round2_sweep.sh build
#   on the box:     gpurun --timeout 900 -- 'tools/round2_sweep.sh run > gpurun_out/round2_sweep.txt 2>&1'
set -uo pipefail
root=$(cd "$(dirname "$0")/.." && pwd)
cd "$root"
case "${1:-}" in
build)
  rm -rf build/variants
  tools/build_variant.sh base
  tools/build_variant.sh fused -DRT_FUSE_NEXT_HIT=1          # shadow test + mirror ray's nearest hit in one loop
  tools/build_variant.sh nopeel -DRT_PEEL_PRIMARY=0          # half the code: launch ramp on small frames
  tools/build_variant.sh fused_mb6 -DRT_FUSE_NEXT_HIT=1 -DRT_MIN_BLOCKS=6   # 80 registers, 6 blocks per SM
  tools/build_variant.sh fused_pairs -DRT_FUSE_NEXT_HIT=1 -DRT_FUSE_PAIRS=1 # two spheres per trip of the shared loop
  ;;
run)
  for v in fused fused_mb6 fused_pairs; do
    echo "== parity with $v"
    RT_B200_LIBRARY=$root/build/variants/$v/librt_b200.so timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -4
  done
  for wl in c1 c2 c4; do
    echo "== workload $wl"
    tools/sweep.sh --workload $wl --steps 100 --warmup 5
  done
  ;;
*) echo "usage: $0 build|run"; exit 2 ;;
esac
