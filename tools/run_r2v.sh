#!/bin/bash
set -x
mkdir -p gpurun_out
L=gpurun_out/r2v_cv.log
: > $L
timeout 300 python tools/devbench_cv2.py >> $L 2>&1
DEVBENCH_PROF=3 timeout 300 python tools/devbench_cv2.py >> $L 2>&1
PLSB_CV_TIMELINE=1 timeout 300 python tools/devbench_cv2.py 20000 1 > gpurun_out/r2v_timeline.log 2>&1
timeout 600 python tools/devbench_loo_c4.py 38000 >> $L 2>&1
