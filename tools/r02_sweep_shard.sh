#!/bin/bash
# Ticket-schedule variants for SHORT launches (one rank's eighth of an 8K frame): tools/shard_bench.py per variant.
#   here: tools/r02_sweep_shard.sh build      on the box: tools/r02_sweep_shard.sh run
set -uo pipefail
root=$(cd "$(dirname "$0")/.." && pwd)
cd "$root"
case "${1:-}" in
build)
  rm -rf build/variants
  make -C python_ray_tracer_b200/csrc >/dev/null
  (tools/build_variant.sh base) &
  (tools/build_variant.sh short32 -DRT_SHORT_LAUNCH_ROUNDS=32) &                       # 1/8 of 8K (43 rounds) deals 70 % statically
  (tools/build_variant.sh short32d50 -DRT_SHORT_LAUNCH_ROUNDS=32 -DRT_DYNAMIC_PCT=50) &
  (tools/build_variant.sh chunk4 -DRT_CHUNK4_ROUNDS=16 -DRT_CHUNK2_ROUNDS=6) &          # fours while more than 16 rounds are left
  wait
  (tools/build_variant.sh mb6 -DRT_SMALL_MIN_BLOCKS=6) &
  (tools/build_variant.sh chunk2_12 -DRT_CHUNK2_ROUNDS=12) &
  wait
  ;;
run)
  for lib in build/variants/*/librt_b200.so; do
    echo "== $(basename "$(dirname "$lib")")"
    RT_B200_LIBRARY=$root/$lib python tools/shard_bench.py 1 8
  done
  ;;
esac
