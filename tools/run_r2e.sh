#!/bin/bash
# round 2, session 4, call 2: two-lane CV with the pre-tiled persistent product kernel (cv_gemm.cu); drop-in e2e trace;
# ncu of the new kernels and of the small-data CV kernel (C3)
set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x --durations=8 > gpurun_out/r2e_pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r2e_pytest.log
L=gpurun_out/r2e_cv_ab.log
: > $L
PLSB_CV_LANES=1 timeout 300 python tools/devbench_cv2.py >> $L 2>&1
timeout 300 python tools/devbench_cv2.py >> $L 2>&1
PLSB_CV_PRIO=0 timeout 300 python tools/devbench_cv2.py >> $L 2>&1
PLSB_FOLD_BATCH=1792 timeout 300 python tools/devbench_cv2.py >> $L 2>&1
PLSB_FOLD_BATCH=5120 timeout 300 python tools/devbench_cv2.py >> $L 2>&1
nproc > gpurun_out/r2e_host.txt; free -g >> gpurun_out/r2e_host.txt; lscpu | head -25 >> gpurun_out/r2e_host.txt
PLSB_TRACE=1 timeout 600 tests/cpp/api_bench fit 1000000 2048 8 32 2 > gpurun_out/r2e_apibench_fit.log 2>&1
PLSB_TRACE=1 PLSB_PACK_THREADS=32 timeout 600 tests/cpp/api_bench fit 1000000 2048 8 32 2 > gpurun_out/r2e_apibench_fit_t32.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"cv_gemm_kernel|cv_fold_fused" --launch-skip 300 --launch-count 4 -o gpurun_out/r2e_cv -f python tools/devbench_cv2.py 7000 1 > gpurun_out/r2e_ncu_cv.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"cv_small_kernel" --launch-skip 4 --launch-count 1 -o gpurun_out/r2e_cvsmall -f python tools/devbench_cv.py 256 > gpurun_out/r2e_ncu_cvsmall.log 2>&1
ls -la gpurun_out | grep r2e
