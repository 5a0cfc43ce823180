#!/bin/bash
cd "${GRAFT_REPO_ROOT:-.}"; mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_fullsize.py tests/test_gpu_reference_dropin.py tests/test_gpu_parity.py tests/test_zz_no_alloc.py -m gpu -x -q 2>&1 | tail -5
timeout 300 python tools/bench_stages.py --reps 20 --only pl_low_cfg2_w12,pl_hist_fused,pl_hist_separate,pl_hist_fused_co,pl_hist_fused_city,pl_hist_separate_city > gpurun_out/r02_fused_hist.jsonl 2> gpurun_out/r02_fused_hist.err; echo rc=$?
python - <<'PY'
import json
for l in open('gpurun_out/r02_fused_hist.jsonl'):
    try: d=json.loads(l)
    except Exception: continue
    print('%-44s %8.2f us (min %8.2f)'%(d['stage'], d['us_median'], d['us_min']))
PY
tail -3 gpurun_out/r02_fused_hist.err
for m in coresident coresident_hist_fused; do
timeout 300 python bench.py --steps 50 --warmup 5 --no-aux --no-cpu-baseline --overlap $m 2> gpurun_out/r02_fh_$m.err | tee gpurun_out/r02_fh_$m.json | python -c "
import sys,json
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$m', round(d['ms_per_step'],5), round(d['value']/1e9,1), d['parity'], d['e2e']['ms_per_step'])"
tail -2 gpurun_out/r02_fh_$m.err
done
