#!/bin/bash
# One script for the GPU-side checks of a tree (it replaces the per-session run_r*.sh scratch files of rounds 1-2).
# Run it ON the GPU box, from the repo root:   tools/gpu_check.sh <tag> [N] [what...]
#   tag   prefix of the files written under gpurun_out/ (e.g. r3x)
#   N     GPUs (default 1); N > 1 runs bench.py under torchrun, one rank per GPU
#   what  any of: tests multi smoke bench reference midcv shape   (default: tests smoke bench)
#   BENCH_FLAGS (environment): extra bench.py flags, e.g. BENCH_FLAGS=--e2e-cpp-multi
# Every step runs under its own timeout and writes <tag>_<step>[_nN].log / .json plus its exit code.
TAG=${1:?usage: tools/gpu_check.sh <tag> [N] [what...]}; N=${2:-1}; shift; shift
WHAT=${*:-tests smoke bench}
mkdir -p gpurun_out
O=gpurun_out/$TAG
for w in $WHAT; do
  case $w in
    tests)     timeout 1500 python -m pytest tests -m gpu -q -x --durations=5 > ${O}_pytest_gpu_n$N.log 2>&1; echo "pytest rc=$?" | tee -a ${O}_pytest_gpu_n$N.log ;;
    multi)     timeout 900 python -m pytest tests/test_gpu_multi.py -m gpu -q -x -rs > ${O}_pytest_multi_n$N.log 2>&1; echo "multi rc=$?" | tee -a ${O}_pytest_multi_n$N.log ;;
    smoke)     timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > ${O}_smoke.log 2>&1; echo "smoke rc=$?" | tee -a ${O}_smoke.log ;;
    bench)     if [ "$N" -gt 1 ]; then
                 timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 \
                     bench.py --gpus $N --steps 20 --warmup 3 $BENCH_FLAGS > ${O}_bench_n$N.json 2> ${O}_bench_n$N.err
               else
                 timeout 1500 python bench.py --steps 20 --warmup 3 > ${O}_bench_n1.json 2> ${O}_bench_n1.err
               fi; echo "bench rc=$?" ;;
    reference) timeout 1500 python bench.py --impl reference --steps 3 --warmup 1 > ${O}_bench_reference.json 2> ${O}_bench_reference.err; echo "reference rc=$?" ;;
    midcv)     timeout 300 python tools/devbench_midcv.py > ${O}_midcv.log 2>&1; echo "midcv rc=$?" ;;
    shape)     timeout 400 python tools/devbench_shape.py 3 > ${O}_shape.log 2>&1; echo "shape rc=$?" ;;
    *)         echo "unknown step $w" ;;
  esac
done
ls -la gpurun_out | grep "$TAG"
