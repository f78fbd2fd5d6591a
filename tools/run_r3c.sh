#!/bin/bash
set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_gpu_cv.py tests/test_gpu_fit.py -m gpu -q -x > gpurun_out/r3c_pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r3c_pytest.log
timeout 600 python tools/devbench_misc.py > gpurun_out/r3c_misc.log 2>&1
