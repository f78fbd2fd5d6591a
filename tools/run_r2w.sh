#!/bin/bash
set -x
mkdir -p gpurun_out
L=gpurun_out/r2w_cv.log
: > $L
DEVBENCH_PROF=3 timeout 300 python tools/devbench_cv2.py >> $L 2>&1
DEVBENCH_PROF=3 PLSB_CV_SERIAL=1 timeout 300 python tools/devbench_cv2.py >> $L 2>&1
PLSB_CV_RING=4 timeout 300 python tools/devbench_cv2.py >> $L 2>&1
