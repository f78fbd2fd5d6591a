#!/bin/bash
set -x
mkdir -p gpurun_out
PLSB_CV_TIMELINE=1 timeout 300 python tools/devbench_cv2.py 20000 1 > gpurun_out/r2t_timeline.log 2>&1
