#!/bin/bash
TAG=${1:-r02e}
cd "${GRAFT_REPO_ROOT:-.}"; mkdir -p gpurun_out
timeout 300 python tools/bench_stages.py --reps 12 --only pl_argmax_cfg2,pl_low_cfg2,hist_u8_c13 > gpurun_out/${TAG}_stages.jsonl 2> gpurun_out/${TAG}_stages.err; echo "stages rc=$?"; cut -c1-140 gpurun_out/${TAG}_stages.jsonl
timeout 200 python tools/prof_stage.py dist 2 > gpurun_out/${TAG}_prof_plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:"distance_stream" -c 2 -o gpurun_out/${TAG}_stream python tools/prof_stage.py dist 2 > gpurun_out/${TAG}_ncu.log 2>&1; echo "ncu rc=$?"; tail -2 gpurun_out/${TAG}_ncu.log
