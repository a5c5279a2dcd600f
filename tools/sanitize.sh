#!/bin/bash
# compute-sanitizer over small renders of every kernel family (GPU box only; ONE tool per gpurun call --
# B200_PROFILING.md): 160x90 README frame through the kernel-parameter path, the staged-table path
# (RT_FLAG_GENERIC_KERNEL) in both block shapes' code, the sphere tree, explicit rays, both policies.
#   gpurun --timeout 900 -- 'tools/sanitize.sh memcheck  > gpurun_out/sanitize_memcheck.txt 2>&1'
#   gpurun --timeout 900 -- 'tools/sanitize.sh racecheck > gpurun_out/sanitize_racecheck.txt 2>&1'
set -uo pipefail
tool=${1:-memcheck}
root=$(cd "$(dirname "$0")/.." && pwd)
cd "$root"
timeout 800 compute-sanitizer --tool "$tool" --error-exitcode 3 python tools/sanitize_workload.py
echo "compute-sanitizer --tool $tool exit code: $?"
