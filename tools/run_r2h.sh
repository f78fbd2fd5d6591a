#!/bin/bash
# round 2, session 4, call 5: L2 bulk prefetch in the fused fold kernel (A/B), M = 8 fold kernel at 80 registers,
# small-call overheads (trace), full bench
set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x --durations=8 > gpurun_out/r2h_pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r2h_pytest.log
L=gpurun_out/r2h_cv_ab.log
: > $L
PLSB_CV_LANES=1 timeout 300 python tools/devbench_cv2.py >> $L 2>&1
PLSB_CV_LANES=1 PLSB_CV_PREFETCH=1 timeout 300 python tools/devbench_cv2.py >> $L 2>&1
timeout 300 python tools/devbench_cv2.py >> $L 2>&1
PLSB_CV_PREFETCH=0 timeout 300 python tools/devbench_cv2.py >> $L 2>&1
DEVBENCH_PROF=3 timeout 300 python tools/devbench_cv2.py >> $L 2>&1
DEVBENCH_PROF=3 PLSB_CV_SERIAL=1 timeout 300 python tools/devbench_cv2.py >> $L 2>&1
timeout 300 python tools/trace_c3.py > gpurun_out/r2h_trace_c3.log 2>&1
timeout 1500 python bench.py --steps 5 --warmup 3 --no-cpu > gpurun_out/r2h_bench.log 2>&1
ls -la gpurun_out | grep r2h
