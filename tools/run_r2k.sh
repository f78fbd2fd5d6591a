#!/bin/bash
# ring depth A/B for the fused fold kernel (16 KB slots), tests
set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_gpu_cv.py tests/test_gpu_fullsize.py -m gpu -q -x > gpurun_out/r2k_pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r2k_pytest.log
L=gpurun_out/r2k_cv_ab.log
: > $L
for ns in 4 5 6 3; do
  PLSB_CV_RING=$ns DEVBENCH_PROF=3 timeout 300 python tools/devbench_cv2.py >> $L 2>&1
done
PLSB_CV_RING=5 DEVBENCH_PROF=3 PLSB_CV_SERIAL=1 timeout 300 python tools/devbench_cv2.py >> $L 2>&1
