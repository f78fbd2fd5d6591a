#!/bin/bash
# round 2, session 4, call 1: GPU suite with the two-lane CV engine + persistent GEMM, A/B of the CV schedules at C5,
# ncu --set full of the persistent TMA GEMM and the fold kernel
set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x --durations=8 > gpurun_out/r2d_pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r2d_pytest.log
L=gpurun_out/r2d_cv_ab.log
: > $L
PLSB_CV_LANES=1 timeout 300 python tools/devbench_cv2.py >> $L 2>&1
timeout 300 python tools/devbench_cv2.py >> $L 2>&1
PLSB_CV_PRIO=0 timeout 300 python tools/devbench_cv2.py >> $L 2>&1
PLSB_GEMM_PERSIST=0 timeout 300 python tools/devbench_cv2.py >> $L 2>&1
PLSB_GEMM_SELF=0 timeout 300 python tools/devbench_cv2.py >> $L 2>&1
PLSB_CV_LANES=1 PLSB_GEMM_PERSIST=1 PLSB_GEMM_SELF=1 timeout 300 python tools/devbench_cv2.py >> $L 2>&1
PLSB_CV_LANES=1 PLSB_GEMM_PERSIST=1 timeout 300 python tools/devbench_cv2.py >> $L 2>&1
PLSB_FOLD_BATCH=1792 timeout 300 python tools/devbench_cv2.py >> $L 2>&1
PLSB_FOLD_BATCH=5120 timeout 300 python tools/devbench_cv2.py >> $L 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"dmma_tma|cv_fold_fused" --launch-skip 40 --launch-count 4 -o gpurun_out/r2d_cv -f python tools/devbench_cv2.py 7000 1 > gpurun_out/r2d_ncu_cv.log 2>&1
ls -la gpurun_out | grep r2d
