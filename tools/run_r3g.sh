#!/bin/bash
set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_gpu_cv.py -m gpu -q -x > gpurun_out/r3g_pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r3g_pytest.log
