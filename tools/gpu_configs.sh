#!/bin/bash
# every BASELINE.json configuration through bench.py on one GPU.  usage: bash tools/gpu_configs.sh TAG
TAG=${1:-r02cfg}
cd "${GRAFT_REPO_ROOT:-.}"; mkdir -p gpurun_out
for c in cfg1 cfg3 cfg4 cfg5 cfg2; do
  timeout 600 python bench.py --config $c > gpurun_out/${TAG}_${c}.json 2> gpurun_out/${TAG}_${c}.err; echo "$c rc=$?"; tail -c 500 gpurun_out/${TAG}_${c}.err; cut -c1-260 gpurun_out/${TAG}_${c}.json
done
for c in cfg1 cfg3 cfg4 cfg5 cfg2; do
  timeout 900 python bench.py --impl reference --config $c > gpurun_out/${TAG}_${c}_ref.json 2> gpurun_out/${TAG}_${c}_ref.err; echo "$c ref rc=$?"; cut -c1-200 gpurun_out/${TAG}_${c}_ref.json
done
