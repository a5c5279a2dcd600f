#!/bin/bash
# Run bench.py once per built variant (build/variants/*) and print one summary line each.
# usage: tools/sweep.sh [bench args...]   (on the GPU box)
root=$(cd "$(dirname "$0")/.." && pwd)
cd "$root"
for lib in build/variants/*/librt_b200.so; do
  name=$(basename "$(dirname "$lib")")
  RT_B200_LIBRARY=$root/$lib python bench.py --no-cpu-baseline --no-c4 --no-c3 "$@" 2>&1 | python -c "
import sys, json
name = sys.argv[1]
for line in sys.stdin:
    line = line.strip()
    if line.startswith('{'):
        d = json.loads(line)
        print(f\"{name:14s} kernel_ms={d['ms_per_step']:.4f} value={d['value']:.0f} e2e_ms={d['e2e']['ms_per_step']:.3f} planes_ms={d['e2e'].get('linear_planes_f32', {}).get('ms_per_step', 0):.3f} fp64frac={d.get('roofline', {}).get('frac') or 0:.3f} clocks={d.get('clocks', {}).get('sm_mhz')}\")
    elif line:
        print(name, line[:200])
" "$name"
done
