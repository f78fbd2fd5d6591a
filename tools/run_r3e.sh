#!/bin/bash
set -x
mkdir -p gpurun_out
L=gpurun_out/r3e_cv.log
: > $L
for pf in 0 2 4 8; do
  PLSB_CV_RING_PF=$pf DEVBENCH_PROF=3 timeout 300 python tools/devbench_cv2.py >> $L 2>&1
done
PLSB_CV_RING_PF=4 DEVBENCH_PROF=3 PLSB_CV_SERIAL=1 timeout 300 python tools/devbench_cv2.py >> $L 2>&1
PLSB_CV_RING_PF=4 timeout 600 python tools/devbench_loo_c4.py 38000 >> $L 2>&1
