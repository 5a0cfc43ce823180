#!/bin/bash
# sweep of the co-resident launch forms on the full bench step. usage: bash tools/gpu_sweep_step.sh FORM "a,b,c,d a,b,c,d ..."
cd "${GRAFT_REPO_ROOT:-.}"; mkdir -p gpurun_out; FORM=${1:-coresident_prelabel}; shift
for v in $@; do
timeout 200 python bench.py --steps 50 --warmup 5 --no-aux --no-cpu-baseline --no-parity --overlap $FORM --coresident $v 2>/dev/null | python -c "
import sys,json
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$FORM $v', round(d['ms_per_step'],5), round(d['value']/1e9,1), flush=True)"
done | tee -a gpurun_out/r02_sweep_prelabel.txt
