#!/bin/bash
# multi-GPU pass: the 2-GPU peer-exchange test, then bench.py under torchrun at N ranks.  usage: bash tools/gpu_multi.sh TAG N [steps]
TAG=${1:-r02m}; N=${2:-2}; STEPS=${3:-50}
cd "${GRAFT_REPO_ROOT:-.}"; mkdir -p gpurun_out
nvidia-smi --query-gpu=index,name --format=csv,noheader | head -8
nvidia-smi topo -m > gpurun_out/${TAG}_topo.txt 2>&1
timeout 600 python -m pytest tests/test_gpu_peer_exchange.py -m gpu -q --timeout 500 > gpurun_out/${TAG}_peer_test.log 2>&1; echo "peer test rc=$?"; tail -3 gpurun_out/${TAG}_peer_test.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus $N --steps $STEPS --warmup 5 > gpurun_out/${TAG}_bench_n${N}.json 2> gpurun_out/${TAG}_bench_n${N}.err; echo "bench rc=$?"; tail -c 1200 gpurun_out/${TAG}_bench_n${N}.err; cut -c1-400 gpurun_out/${TAG}_bench_n${N}.json
