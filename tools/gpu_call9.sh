#!/bin/bash
TAG=${1:-r02j}
cd "${GRAFT_REPO_ROOT:-.}"; mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -q --timeout 400 -k "pseudo or fullsize or cfg2" > gpurun_out/${TAG}_tests.log 2>&1; echo "tests rc=$?"; tail -3 gpurun_out/${TAG}_tests.log
timeout 500 python tools/bench_stages.py --reps 12 --only pl_low_cfg2,pl_rows2,pl_rows2_c3w2,pl_rows2_c4w2,pl_low_city,pl_rows2_city,ov_15_6_c3w2_p,ov_r2_12_6_c3w2,ov_r2_12_4_c3w2,ov_r2_12_4_c3w3,ov_r2_12_4_c4w2,ov_r2_12_8_c3w2,ov_r2_12_16_c3w2 > gpurun_out/${TAG}_stages.jsonl 2> gpurun_out/${TAG}_stages.err; echo "stages rc=$?"; tail -2 gpurun_out/${TAG}_stages.err
