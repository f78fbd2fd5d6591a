#!/bin/bash
set -x
mkdir -p gpurun_out
for w in 0 2 3; do
PLSB_CV_CTA_TRACE=40 PLSB_CV_CTA_TRACE_WHICH=$w PLSB_CV_CTA_TRACE_FILE=gpurun_out/r2u_cta_$w.csv timeout 300 python tools/devbench_cv2.py 20000 1 > gpurun_out/r2u_$w.log 2>&1
python tools/cta_trace_stats.py gpurun_out/r2u_cta_$w.csv > gpurun_out/r2u_stats_$w.log 2>&1
done
