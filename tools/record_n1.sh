#!/bin/bash
# One-GPU records of a round (run on the GPU box): GPU suite, smoke, the bench line as the driver runs it, the reference
# arm, C1, the turntable, the ncu launch list of the bench command and one full capture of the headline kernel.
# usage: tools/record_n1.sh <out dir under gpurun_out/>
set -u
cd "$(dirname "$0")/.."
out=gpurun_out/${1:-record}
mkdir -p "$out"
python -m pytest tests -x -q -m gpu > "$out/gputest_n1.txt" 2>&1; echo "rc=$?" >> "$out/gputest_n1.txt"
python __graft_entry__.py smoke > "$out/smoke.txt" 2>&1; echo "rc=$?" >> "$out/smoke.txt"
python bench.py > "$out/bench_n1.json" 2> "$out/bench_n1.err"
python bench.py --impl reference --steps 3 --warmup 1 > "$out/reference_arm_n1.json" 2> "$out/reference_arm_n1.err"
python bench.py --workload c1 --no-cpu-baseline --no-c4 --no-c3 --steps 200 --warmup 5 > "$out/bench_c1.json" 2> "$out/bench_c1.err"
python -m python_ray_tracer_b200.turntable --frames 256 --out /tmp/tt_dev > "$out/turntable_c5.txt" 2>&1
python -m python_ray_tracer_b200.turntable --frames 256 --out /tmp/tt_host --encoder host >> "$out/turntable_c5.txt" 2>&1
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file "$out/launches_bench_first400.csv" \
    python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-c4 --no-c3 > "$out/ncu_launches.log" 2>&1
ncu --set full --clock-control none --import-source on -k regex:render_rows_small_kernel -s 4 -c 2 -f -o "$out/small3" \
    python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-c4 --no-c3 > "$out/ncu_full.log" 2>&1
tail -n 3 "$out/gputest_n1.txt" "$out/smoke.txt"
