#!/bin/bash
# ncu launch list of a short bench.py run (one ncu command, after the same command line exited 0 without it)
TAG=${1:-r02b}
cd "${GRAFT_REPO_ROOT:-.}"; mkdir -p gpurun_out
B="python bench.py --steps 3 --warmup 3 --no-aux --no-cpu-baseline --no-parity --overlap coresident_prelabel"
$B > gpurun_out/${TAG}_ncu_plain_bench.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/${TAG}_launches_bench.csv $B > gpurun_out/${TAG}_ncu_bench.log 2>&1; echo "launch list rc=$?"
