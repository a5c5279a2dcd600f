#!/bin/bash
# Config 4 (8K frame over N ranks) with both exchange modes.  usage: tools/c4_scaling.sh N [steps]
n=$1; steps=${2:-30}
cd "$(dirname "$0")/.."
port=29600
for ex in peer nccl; do
  port=$((port+1))
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $port \
    bench.py --gpus $n --steps $steps --warmup 5 --workload c4 --exchange $ex 2>&1 | tail -n 3 > gpurun_out/c4_n${n}_$ex.log
  python - "$ex" "$n" <<'PY'
import json, sys
ex, n = sys.argv[1], sys.argv[2]
for l in open(f"gpurun_out/c4_n{n}_{ex}.log"):
    if l.startswith("{"):
        d = json.loads(l)
        print(f"c4 n={n} {ex}: ms/frame={d['ms_per_step']:.4f} value={d['value']:.0f} e2e_ms={d['e2e']['ms_per_step']:.3f}")
    elif l.strip():
        print(l[:300].rstrip())
PY
done
