#!/bin/bash
cd "${GRAFT_REPO_ROOT:-.}"; mkdir -p gpurun_out
timeout 200 python -m pytest tests -m gpu -x -q -k "pipeline or Pipeline or host" 2>&1 | tail -3
timeout 200 python bench.py --steps 20 --warmup 3 --no-aux --no-cpu-baseline 2>gpurun_out/pipe.err | python -c "
import sys,json
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(round(d['ms_per_step'],5), d['e2e']['ms_per_step'], d['e2e']['frac_of_bare_copy_ceiling'], d['e2e']['value'])" || tail -5 gpurun_out/pipe.err
