#!/bin/bash
TAG=${1:-r02i}
cd "${GRAFT_REPO_ROOT:-.}"; mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q --timeout 400 > gpurun_out/${TAG}_tests.log 2>&1; echo "tests rc=$?"; tail -3 gpurun_out/${TAG}_tests.log
timeout 300 python tools/bench_stages.py --reps 12 --only pl_low_cfg2,pl_c3w2,pl_argmax_cfg2,serial_cfg2,ov_15_6_c3w2_p,pl_low_city > gpurun_out/${TAG}_stages.jsonl 2> gpurun_out/${TAG}_stages.err; echo "stages rc=$?"; cut -c1-120 gpurun_out/${TAG}_stages.jsonl
timeout 500 python bench.py --steps 50 --warmup 5 > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err; echo "bench rc=$?"; tail -c 300 gpurun_out/${TAG}_bench.err; cut -c1-300 gpurun_out/${TAG}_bench.json
timeout 500 python bench.py --config cfg4 --no-cpu-baseline > gpurun_out/${TAG}_cfg4.json 2> gpurun_out/${TAG}_cfg4.err; echo "cfg4 rc=$?"; cut -c1-300 gpurun_out/${TAG}_cfg4.json
