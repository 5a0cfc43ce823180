#!/bin/bash
TAG=${1:-r02g}
cd "${GRAFT_REPO_ROOT:-.}"; mkdir -p gpurun_out
timeout 600 python tools/bench_stages.py --reps 10 --only dist_stream_shapes,dist_stream_shapes_k256,loss_cfg2 > gpurun_out/${TAG}_stages.jsonl 2> gpurun_out/${TAG}_stages.err; echo "stages rc=$?"; tail -3 gpurun_out/${TAG}_stages.err
