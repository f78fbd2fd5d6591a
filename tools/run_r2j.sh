#!/bin/bash
# phase timestamps of one fused fold CTA: next to the product kernel (two lanes) and alone (serial)
set -x
mkdir -p gpurun_out
L=gpurun_out/r2j_trace.log
: > $L
for a in 4 32 60; do
  PLSB_CV_TRACE=$a timeout 300 python tools/devbench_cv2.py 7000 1 >> $L 2>&1
  PLSB_CV_TRACE=$a PLSB_CV_SERIAL=1 timeout 300 python tools/devbench_cv2.py 7000 1 >> $L 2>&1
done
PLSB_CV_TRACE=32 PLSB_CV_RING=0 timeout 300 python tools/devbench_cv2.py 7000 1 >> $L 2>&1
PLSB_CV_TRACE=32 PLSB_CV_RING=0 PLSB_CV_SERIAL=1 timeout 300 python tools/devbench_cv2.py 7000 1 >> $L 2>&1
