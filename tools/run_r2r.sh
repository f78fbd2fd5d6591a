#!/bin/bash
set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_gpu_cv.py tests/test_gpu_fullsize.py -m gpu -q -x > gpurun_out/r2r_pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r2r_pytest.log
L=gpurun_out/r2r_loo_c4.log
: > $L
DEVBENCH_PROF=3 timeout 600 python tools/devbench_loo_c4.py 38000 >> $L 2>&1
DEVBENCH_PROF=3 PLSB_CV_SERIAL=1 timeout 600 python tools/devbench_loo_c4.py 38000 >> $L 2>&1
PLSB_CV_M8_MINB=4 DEVBENCH_PROF=3 timeout 600 python tools/devbench_loo_c4.py 38000 >> $L 2>&1
PLSB_CV_M8_MINB=4 DEVBENCH_PROF=3 PLSB_CV_SERIAL=1 timeout 600 python tools/devbench_loo_c4.py 38000 >> $L 2>&1
PLSB_CV_RING=4 DEVBENCH_PROF=3 PLSB_CV_SERIAL=1 timeout 600 python tools/devbench_loo_c4.py 38000 >> $L 2>&1
