#!/bin/bash
# Device PNG encoder once per built variant (build/variants/*): band / pack kernel times (tools/png_device_bench.py).
cd "$(dirname "$0")/.."
for lib in build/variants/*/librt_b200.so; do
  name=$(basename $(dirname $lib))
  for size in "1920 1080" "7680 4320"; do
    echo "== $name $size"
    RT_B200_LIBRARY=$PWD/$lib python tools/png_device_bench.py $size 20 2>&1 | grep -E "band kernel|rows per band|Error|error" | head -4
  done
done
