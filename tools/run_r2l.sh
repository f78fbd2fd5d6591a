#!/bin/bash
# how many SMs the product kernel should leave entirely to the fold kernel (A/B), tests
set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_gpu_cv.py tests/test_gpu_fullsize.py -m gpu -q -x > gpurun_out/r2l_pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r2l_pytest.log
L=gpurun_out/r2l_cv_ab.log
: > $L
for g in 148 136 128 120 112; do
  PLSB_CVG_CTAS=$g DEVBENCH_PROF=3 timeout 300 python tools/devbench_cv2.py >> $L 2>&1
done
timeout 300 python tools/devbench_cv2.py >> $L 2>&1
