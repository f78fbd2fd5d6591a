#!/bin/bash
set -x
mkdir -p gpurun_out
timeout 300 python tools/trace_c3.py > gpurun_out/r2o_trace_c3.log 2>&1
PLSB_D2H_KERNEL=0 timeout 300 python tools/trace_c3.py > gpurun_out/r2o_trace_c3_dma.log 2>&1
bash tools/profile_r2.sh
