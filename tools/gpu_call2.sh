#!/bin/bash
TAG=${1:-r02b}
cd "${GRAFT_REPO_ROOT:-.}"; mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_parity.py -m gpu -q -k "finalize_scatter or class_means" --timeout 300 > gpurun_out/${TAG}_tests.log 2>&1; echo "tests rc=$?"; tail -3 gpurun_out/${TAG}_tests.log
timeout 400 python tools/bench_stages.py --reps 20 --only serial_cfg2,overlap_cfg2_12_4,overlap_cfg2_10_6,overlap_cfg2_12_6,overlap_cfg2_8_8,overlap_cfg2_16_4,pl_low_cfg2_w12,pl_low_cfg2_w10,sums_cfg2_nc6,sums_cfg2_nc4 > gpurun_out/${TAG}_stages.jsonl 2> gpurun_out/${TAG}_stages.err; echo "stages rc=$?"; cat gpurun_out/${TAG}_stages.jsonl | cut -c1-150
timeout 500 python bench.py --steps 30 --warmup 3 > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err; echo "bench rc=$?"; tail -c 800 gpurun_out/${TAG}_bench.err; cut -c1-1500 gpurun_out/${TAG}_bench.json
timeout 200 python tools/prof_stage.py distce 2 > gpurun_out/${TAG}_prof_plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:"distance_tlp|ce_rce_fwd" -c 6 -o gpurun_out/${TAG}_dist python tools/prof_stage.py distce 2 > gpurun_out/${TAG}_ncu.log 2>&1; echo "ncu rc=$?"; tail -3 gpurun_out/${TAG}_ncu.log
