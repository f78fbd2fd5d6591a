#!/bin/bash
set -x
mkdir -p gpurun_out
PLSB_CV_CTA_TRACE=32 PLSB_CV_CTA_TRACE_FILE=gpurun_out/r2s_cta_overlap.csv timeout 300 python tools/devbench_cv2.py 20000 1 > gpurun_out/r2s_a.log 2>&1
PLSB_CV_SERIAL=1 PLSB_CV_CTA_TRACE=32 PLSB_CV_CTA_TRACE_FILE=gpurun_out/r2s_cta_serial.csv timeout 300 python tools/devbench_cv2.py 20000 1 > gpurun_out/r2s_b.log 2>&1
python tools/cta_trace_stats.py gpurun_out/r2s_cta_overlap.csv > gpurun_out/r2s_stats.log 2>&1
python tools/cta_trace_stats.py gpurun_out/r2s_cta_serial.csv >> gpurun_out/r2s_stats.log 2>&1
