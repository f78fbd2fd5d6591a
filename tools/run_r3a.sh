#!/bin/bash
set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_gpu_cv.py tests/test_gpu_fit.py -m gpu -q -x > gpurun_out/r3a_pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r3a_pytest.log
L=gpurun_out/r3a_cv.log
: > $L
timeout 300 python tools/devbench_cv2.py >> $L 2>&1
timeout 300 python tools/devbench_cv2.py 5000 3 >> $L 2>&1
timeout 300 python tools/devbench_cv2.py 2500 3 >> $L 2>&1
timeout 600 python tools/devbench.py 250000 8192 8 8 > gpurun_out/r3a_wideK.log 2>&1
