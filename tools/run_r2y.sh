#!/bin/bash
set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_gpu_fit.py -m gpu -q -x > gpurun_out/r2y_pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r2y_pytest.log
timeout 600 python tools/devbench.py 250000 8192 8 8 > gpurun_out/r2y_wideK.log 2>&1
timeout 600 python tools/devbench.py 400000 5000 8 8 > gpurun_out/r2y_wideK5000.log 2>&1
PLSB_PASS_WIDE_V1=1 timeout 600 python tools/devbench.py 400000 5000 8 8 > gpurun_out/r2y_wideK5000_v1.log 2>&1
