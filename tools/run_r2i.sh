#!/bin/bash
# round 2, session 4, call 6: TMA-fed ring for the fold kernel's P/R stream + 2-stage product kernel (A/B), download trace
set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x --durations=8 > gpurun_out/r2i_pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r2i_pytest.log
L=gpurun_out/r2i_cv_ab.log
: > $L
timeout 300 python tools/devbench_cv2.py >> $L 2>&1
PLSB_CV_RING=0 timeout 300 python tools/devbench_cv2.py >> $L 2>&1
PLSB_CV_LANES=1 timeout 300 python tools/devbench_cv2.py >> $L 2>&1
PLSB_CV_LANES=1 PLSB_CV_RING=0 timeout 300 python tools/devbench_cv2.py >> $L 2>&1
DEVBENCH_PROF=3 timeout 300 python tools/devbench_cv2.py >> $L 2>&1
DEVBENCH_PROF=3 PLSB_CV_SERIAL=1 timeout 300 python tools/devbench_cv2.py >> $L 2>&1
DEVBENCH_PROF=3 PLSB_CV_SERIAL=1 PLSB_CVG_STAGES=3 timeout 300 python tools/devbench_cv2.py >> $L 2>&1
PLSB_CV_PRIO=0 timeout 300 python tools/devbench_cv2.py >> $L 2>&1
timeout 300 python tools/trace_c3.py > gpurun_out/r2i_trace_c3.log 2>&1
ls -la gpurun_out | grep r2i
