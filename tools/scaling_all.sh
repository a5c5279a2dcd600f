#!/bin/bash
# N = 1, 2, 4, 8 for configs 2 (frame batch, weak) and 4 (one 8K frame over N ranks, strong), one
# 8-GPU box.  usage: tools/scaling_all.sh [steps]   -> gpurun_out/scaling_<workload>_n<N>.json
steps=${1:-100}
cd "$(dirname "$0")/.."
port=29700
for w in c2 c4; do
  for n in 1 2 4 8; do
    port=$((port+1))
    if [ "$n" = 1 ]; then
      python bench.py --gpus 1 --steps $steps --warmup 5 --workload $w --no-cpu-baseline 2>/dev/null | tail -n 1 > gpurun_out/scaling_${w}_n$n.json
    else
      python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $port \
        bench.py --gpus $n --steps $steps --warmup 5 --workload $w 2>/dev/null | tail -n 1 > gpurun_out/scaling_${w}_n$n.json
    fi
    python - "$w" "$n" <<'PY'
import json, sys
w, n = sys.argv[1], sys.argv[2]
try:
    d = json.loads(open(f"gpurun_out/scaling_{w}_n{n}.json").read())
    print(f"{w} n={n}: ms/step={d['ms_per_step']:.4f} value={d['value']:.0f} {d['unit']} e2e_ms={d['e2e']['ms_per_step']:.3f} e2e_value={d['e2e']['value']:.0f} scaling={d['scaling']}")
except Exception as err:
    print(f"{w} n={n}: no line ({err})")
PY
  done
done
