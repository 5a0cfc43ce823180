#!/bin/bash
# ncu evidence for the judge: (1) launch list of a short bench.py run, (2) --set full of a late (steady-state) launch of every
# kernel of the step.  One gpurun call; each ncu command follows a plain run of the same command line that exited 0.
TAG=${1:-r02}
cd "${GRAFT_REPO_ROOT:-.}"; mkdir -p gpurun_out
B="python bench.py --steps 3 --warmup 3 --no-aux --no-cpu-baseline --no-parity --overlap coresident"
$B > gpurun_out/${TAG}_ncu_plain_bench.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/${TAG}_launches_bench.csv $B > gpurun_out/${TAG}_ncu_bench.log 2>&1; echo "launch list rc=$?"
P="python tools/prof_stage.py all 4"
$P > gpurun_out/${TAG}_ncu_plain_prof.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:"pl_up|proto_accum|proto_label|proto_update|confusion_packed" -s 10 -c 5 -o gpurun_out/${TAG}_step_full $P > gpurun_out/${TAG}_ncu_prof.log 2>&1; echo "full rc=$?"; tail -2 gpurun_out/${TAG}_ncu_prof.log
