#!/bin/bash
TAG=${1:-r02d}
cd "${GRAFT_REPO_ROOT:-.}"; mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q --timeout 400 > gpurun_out/${TAG}_tests.log 2>&1; echo "tests rc=$?"; tail -4 gpurun_out/${TAG}_tests.log
timeout 600 python tools/bench_stages.py --reps 12 --only dist_stream_cfg3,dist_stream_cfg3_4,dist_stream_cfg3_8,dist_stream_k256,dist_tlp_k256_best,ov_15_6_c3w2,ov_15_6_c3w2_p,ov_15_6_c3w2_plf,ov_15_6_c3w3_plf,ov_15_6_c3w4_plf,ov_16_6_c4w2_plf,ov_15_6_c5w2_plf,ov_12_6_c3w2_plf,ov_15_6_c3w1_plf > gpurun_out/${TAG}_stages.jsonl 2> gpurun_out/${TAG}_stages.err; echo "stages rc=$?"; tail -3 gpurun_out/${TAG}_stages.err
timeout 500 python bench.py --steps 30 --warmup 3 > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err; echo "bench rc=$?"; tail -c 800 gpurun_out/${TAG}_bench.err
