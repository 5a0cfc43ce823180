#!/bin/bash
TAG=${1:-r02f}
cd "${GRAFT_REPO_ROOT:-.}"; mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_parity.py -m gpu -q -k "distance" --timeout 300 > gpurun_out/${TAG}_tests.log 2>&1; echo "tests rc=$?"; tail -2 gpurun_out/${TAG}_tests.log
timeout 300 python tools/bench_stages.py --reps 12 --only dist_stream_cfg3,dist_stream_cfg3_4,dist_stream_cfg3_8,dist_stream_k256 > gpurun_out/${TAG}_stages.jsonl 2> gpurun_out/${TAG}_stages.err; echo "stages rc=$?"; cut -c1-140 gpurun_out/${TAG}_stages.jsonl
timeout 200 python tools/prof_stage.py pl 3 > gpurun_out/${TAG}_prof_plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:"pl_up" -s 1 -c 1 -o gpurun_out/${TAG}_pl python tools/prof_stage.py pl 3 > gpurun_out/${TAG}_ncu.log 2>&1; echo "ncu rc=$?"; tail -2 gpurun_out/${TAG}_ncu.log
