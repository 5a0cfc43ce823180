#!/bin/bash
cd "${GRAFT_REPO_ROOT:-.}"; mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_fullsize.py tests/test_gpu_step.py -m gpu -x -q 2>&1 | tail -3
for m in coresident coresident_prelabel auto; do
timeout 300 python bench.py --steps 50 --warmup 5 --no-aux --no-cpu-baseline --overlap $m 2> gpurun_out/r02_pre_$m.err | tee gpurun_out/r02_pre_$m.json | python -c "
import sys,json
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$m', round(d['ms_per_step'],5), round(d['value']/1e9,1), d['config']['step_form_trials_ms'], d['gpu_launches'], d['parity'])"
tail -2 gpurun_out/r02_pre_$m.err
done
