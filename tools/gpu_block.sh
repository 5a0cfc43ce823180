#!/bin/bash
cd "${GRAFT_REPO_ROOT:-.}"; mkdir -p gpurun_out
for extra in "--steps-per-graph 1" "--steps-per-graph 8" "--steps-per-graph 16" "" "--overlap coresident --steps-per-graph 8"; do
timeout 300 python bench.py --steps 50 --warmup 5 --no-aux --no-cpu-baseline $extra 2>gpurun_out/blk.err | python -c "
import sys,json
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('[$extra]', round(d['ms_per_step'],5), round(d['wall_ms_per_step'],5), round(d['value']/1e9,1), d['config']['steps_per_graph'], d['config']['steps_per_graph_trial'], d['parity'].get('hist_exact'), d['config']['miou'], d['gpu_launches'], flush=True)" || tail -5 gpurun_out/blk.err
done | tee gpurun_out/r02_blocks_n1.txt
