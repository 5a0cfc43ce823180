#!/bin/bash
cd "${GRAFT_REPO_ROOT:-.}"; mkdir -p gpurun_out
for extra in "--overlap coresident_prelabel" "--overlap coresident_prelabel --main-streams 2" "--overlap coresident --main-streams 2" "--overlap serial --main-streams 2" "--overlap coresident_prelabel --main-streams 2 --coresident 12,3,2,6" "--overlap coresident_prelabel --main-streams 2 --coresident 16,4,1,6"; do
timeout 200 python bench.py --steps 50 --warmup 5 --no-aux --no-cpu-baseline $extra 2>gpurun_out/ms.err | python -c "
import sys,json
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$extra', round(d['ms_per_step'],5), round(d['value']/1e9,1), d['parity'].get('hist_exact'), d['parity'].get('proto_mean_max_rel_err'), flush=True)" || tail -3 gpurun_out/ms.err
done | tee gpurun_out/r02_main_streams.txt
