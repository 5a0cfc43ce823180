#!/bin/bash
# final single-GPU pass of the round: GPU tests, smoke, the default bench line, the configurations whose step form changed
TAG=${1:-r02fin3}
cd "${GRAFT_REPO_ROOT:-.}"; mkdir -p gpurun_out
timeout 1000 python -m pytest tests -m gpu -x -q --timeout 400 > gpurun_out/${TAG}_tests.log 2>&1; echo "tests rc=$?"; tail -3 gpurun_out/${TAG}_tests.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${TAG}_smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/${TAG}_smoke.log
timeout 600 python bench.py > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err; echo "bench rc=$?"; tail -c 300 gpurun_out/${TAG}_bench.err; cut -c1-300 gpurun_out/${TAG}_bench.json
for c in cfg1 cfg4; do
  timeout 600 python bench.py --config $c > gpurun_out/${TAG}_${c}.json 2> gpurun_out/${TAG}_${c}.err; echo "$c rc=$?"; cut -c1-260 gpurun_out/${TAG}_${c}.json
done
