#!/bin/bash
# round 2, session 4, call 3: carve-out A/B for the two-lane CV, squaring solver in the fold kernels, staged D2H of the
# scores in the drop-in, the new bench legs
set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x --durations=8 > gpurun_out/r2f_pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r2f_pytest.log
L=gpurun_out/r2f_cv_ab.log
: > $L
PLSB_CV_LANES=1 timeout 300 python tools/devbench_cv2.py >> $L 2>&1
timeout 300 python tools/devbench_cv2.py >> $L 2>&1
PLSB_CV_CARVEOUT=0 timeout 300 python tools/devbench_cv2.py >> $L 2>&1
DEVBENCH_PROF=3 timeout 300 python tools/devbench_cv2.py >> $L 2>&1
DEVBENCH_PROF=3 PLSB_CV_SERIAL=1 timeout 300 python tools/devbench_cv2.py >> $L 2>&1
DEVBENCH_PROF=3 PLSB_CV_CARVEOUT=0 timeout 300 python tools/devbench_cv2.py >> $L 2>&1
timeout 300 python tools/devbench_cv.py 256 > gpurun_out/r2f_cv_small.log 2>&1
PLSB_TRACE=1 timeout 600 tests/cpp/api_bench fit 1000000 2048 8 32 3 > gpurun_out/r2f_apibench_fit.log 2>&1
timeout 1200 python bench.py --steps 3 --warmup 3 --no-cpu > gpurun_out/r2f_bench.log 2>&1
ls -la gpurun_out | grep r2f
