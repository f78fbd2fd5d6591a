#!/bin/bash
# round 2, session 4, call 4: column-owner dots in the fused fold kernel (MLP for the lone co-resident CTA), warp-per-fold
# small-data kernel, staged E download
set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x --durations=8 > gpurun_out/r2g_pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r2g_pytest.log
L=gpurun_out/r2g_cv_ab.log
: > $L
PLSB_CV_LANES=1 timeout 300 python tools/devbench_cv2.py >> $L 2>&1
timeout 300 python tools/devbench_cv2.py >> $L 2>&1
DEVBENCH_PROF=3 timeout 300 python tools/devbench_cv2.py >> $L 2>&1
DEVBENCH_PROF=3 PLSB_CV_SERIAL=1 timeout 300 python tools/devbench_cv2.py >> $L 2>&1
timeout 300 python tools/devbench_cv.py 256 > gpurun_out/r2g_cv_small.log 2>&1
PLSB_CV_NO_WARP=1 timeout 300 python tools/devbench_cv.py 256 > gpurun_out/r2g_cv_small_nowarp.log 2>&1
timeout 1200 python bench.py --steps 3 --warmup 3 --no-cpu --no-e2e --no-xtx > gpurun_out/r2g_bench.log 2>&1
ls -la gpurun_out | grep r2g
