#!/bin/bash
set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x --durations=5 > gpurun_out/r2q_pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r2q_pytest.log
timeout 600 python tools/devbench.py 250000 8192 8 8 > gpurun_out/r2q_wideK.log 2>&1
L=gpurun_out/r2q_loo_c4.log
: > $L
DEVBENCH_PROF=3 timeout 600 python tools/devbench_loo_c4.py 38000 >> $L 2>&1
DEVBENCH_PROF=3 PLSB_CV_SERIAL=1 timeout 600 python tools/devbench_loo_c4.py 38000 >> $L 2>&1
PLSB_CV_LANES=1 timeout 600 python tools/devbench_loo_c4.py 38000 >> $L 2>&1
timeout 300 python tools/devbench_cv2.py > gpurun_out/r2q_cv.log 2>&1
