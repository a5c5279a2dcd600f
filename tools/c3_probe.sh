#!/bin/bash
# C3 (1 025 shapes, 4K, depth 8, through the tree) once per built variant (build/variants/*): kernel time and counters.
# usage (on the GPU box): tools/c3_probe.sh [steps]
cd "$(dirname "$0")/.."
for lib in build/variants/*/librt_b200.so; do
  name=$(basename $(dirname $lib))
  RT_B200_LIBRARY=$PWD/$lib python bench.py --workload c3 --no-cpu-baseline --no-c4 --steps ${1:-10} --warmup 3 2>&1 | python -c "
import sys, json
for line in sys.stdin:
    if line.startswith('{'):
        d = json.loads(line); print('$name', round(d['ms_per_step'], 4), d['roofline']['counters'])
    elif 'Error' in line or 'error' in line:
        print('$name', line.strip()[:300])
"
done
