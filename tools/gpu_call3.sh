#!/bin/bash
TAG=${1:-r02c}
cd "${GRAFT_REPO_ROOT:-.}"; mkdir -p gpurun_out
timeout 600 python tools/bench_stages.py --reps 12 --only serial_cfg2,pl_c3w2,ov_15_8_c3w2,ov_15_8_c3w2_p,ov_15_8_c3w2_plf,ov_15_8_c3w3_plf,ov_15_16_c3w2_plf,ov_15_6_c3w2_plf,ov_15_4_c3w2_plf,ov_15_8_c5w2_plf,ov_12_8_c6w2_plf,ov_12_6_plf,ov_12_6_p,overlap_cfg2_12_6 > gpurun_out/${TAG}_stages.jsonl 2> gpurun_out/${TAG}_stages.err; echo "stages rc=$?"; tail -3 gpurun_out/${TAG}_stages.err
timeout 900 python tools/bench_stages.py --reps 8 --only dist_sweep_cfg3,dist_sweep_cfg3_8,dist_sweep_k256 > gpurun_out/${TAG}_dist.jsonl 2> gpurun_out/${TAG}_dist.err; echo "dist rc=$?"; tail -3 gpurun_out/${TAG}_dist.err
