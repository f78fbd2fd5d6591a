#!/bin/bash
# multi-GPU evidence: the multi-GPU tests (sharded fits, fold-sharded CV, cross products across ranks, C++ layer) and
# bench.py at N GPUs (one process per GPU under torchrun).  usage: tools/run_r2_multi.sh <N>
N=${1:-2}
set -x
mkdir -p gpurun_out
nvidia-smi --query-gpu=index,name --format=csv > gpurun_out/r2m_gpus_n$N.txt
nvidia-smi topo -m >> gpurun_out/r2m_gpus_n$N.txt 2>&1
timeout 1500 python -m pytest tests/test_gpu_multi.py -m gpu -q -x -rs --durations=8 > gpurun_out/r2m_pytest_multi_n$N.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r2m_pytest_multi_n$N.log
timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 \
    bench.py --gpus $N --steps 10 --warmup 3 --no-cpu > gpurun_out/r2m_bench_n$N.log 2>&1
echo "bench rc=$?" >> gpurun_out/r2m_bench_n$N.log
ls -la gpurun_out | grep r2m
