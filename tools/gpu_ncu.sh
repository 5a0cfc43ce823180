#!/bin/bash
# ncu evidence, part 2: --set full of a late (steady-state) launch of every kernel of the step.  ONE ncu command per gpurun
# call, after a plain run of the same command line exited 0 (the launch list of bench.py is tools/gpu_launchlist.sh).
TAG=${1:-r02}
cd "${GRAFT_REPO_ROOT:-.}"; mkdir -p gpurun_out
P="python tools/prof_stage.py all 4"
$P > gpurun_out/${TAG}_ncu_plain_prof.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:"pl_up|proto_accum|proto_label|proto_update|confusion_packed" -s 10 -c 5 -o gpurun_out/${TAG}_step_full $P > gpurun_out/${TAG}_ncu_prof.log 2>&1; echo "full rc=$?"; tail -2 gpurun_out/${TAG}_ncu_prof.log
