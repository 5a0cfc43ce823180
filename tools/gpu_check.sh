#!/bin/bash
# One GPU-box pass: parity tests, smoke, per-stage sweeps, a short bench.  usage (through gpurun): bash tools/gpu_check.sh TAG
TAG=${1:-r02}
cd "${GRAFT_REPO_ROOT:-.}"
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.max.sm --format=csv,noheader
timeout 1000 python -m pytest tests -m gpu -q --timeout 400 > gpurun_out/${TAG}_tests.log 2>&1; echo "tests rc=$?"; tail -5 gpurun_out/${TAG}_tests.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${TAG}_smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/${TAG}_smoke.log
timeout 500 python tools/bench_stages.py --reps 20 --only "${STAGES:-serial_cfg2,overlap_cfg2_12_4,overlap_cfg2_8_8,overlap_cfg2_8_4,overlap_cfg2_16_4,pl_low_cfg2,pl_low_cfg2_w12,pl_low_cfg2_w8,sums_cfg2_nc16,sums_cfg2_nc8,sums_cfg2_nc4,proto_cfg2,proto_cfg3,dist_rectify_cfg3,dist_rectify_cfg3_direct,dist_rectify_cfg3_8img,dist_rectify_cfg3_8img_direct,dist_rectify_k256,dist_rectify_k256_direct,loss_cfg2,hist_u8}" > gpurun_out/${TAG}_stages.jsonl 2> gpurun_out/${TAG}_stages.err; echo "stages rc=$?"
timeout 500 python bench.py --steps 30 --warmup 3 > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err; echo "bench rc=$?"; tail -c 800 gpurun_out/${TAG}_bench.err
