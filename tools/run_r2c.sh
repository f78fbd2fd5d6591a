#!/bin/bash
# round 2: GPU suite after the host-layer rewrite, the drop-in's end to end (api_bench), short bench
set -x
mkdir -p gpurun_out
timeout 2400 python -m pytest tests -m gpu -q -x --durations=8 > gpurun_out/r2c_pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r2c_pytest.log
timeout 600 tests/cpp/api_bench fit 1000000 2048 8 32 3 > gpurun_out/r2c_apibench_fit.log 2>&1
PLSB_STAGE_WC=0 timeout 600 tests/cpp/api_bench fit 1000000 2048 8 32 3 > gpurun_out/r2c_apibench_fit_nowc.log 2>&1
PLSB_NO_STAGE=1 timeout 600 tests/cpp/api_bench fit 1000000 2048 8 32 2 > gpurun_out/r2c_apibench_fit_nostage.log 2>&1
PLSB_PACK_THREADS=8 timeout 600 tests/cpp/api_bench fit 1000000 2048 8 32 3 > gpurun_out/r2c_apibench_fit_t8.log 2>&1
timeout 600 tests/cpp/api_bench loo 20000 4096 4 64 1 20260005 > gpurun_out/r2c_apibench_loo.log 2>&1
timeout 1200 python bench.py --steps 5 --warmup 3 --no-cpu > gpurun_out/r2c_bench.log 2>&1
