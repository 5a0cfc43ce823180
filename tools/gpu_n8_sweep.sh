#!/bin/bash
cd "${GRAFT_REPO_ROOT:-.}"; N=${1:-8}
for extra in "--overlap coresident" "--overlap coresident --tail-priority 1" "--overlap coresident_hist_main" "--overlap coresident_hist_main --tail-priority 1"; do
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29544 bench.py --gpus $N --steps 50 --warmup 4 --no-aux --no-cpu-baseline --no-parity $extra 2>/dev/null | python -c "
import sys,json
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$extra', round(d['ms_per_step'],5), round(d['value']/1e9,1), d['config']['bank_vs_sequential_rel'])"
done
