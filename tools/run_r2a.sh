#!/bin/bash
# round 2, first GPU call: full GPU test suite, CV A/B of the GEMM variants, short bench, ncu of the new GEMM + fold kernel
set -x
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,memory.total --format=csv > gpurun_out/r2a_gpu.txt
free -g >> gpurun_out/r2a_gpu.txt; nproc >> gpurun_out/r2a_gpu.txt
timeout 1500 python -m pytest tests -m gpu -q -x --durations=15 > gpurun_out/r2a_pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r2a_pytest.log
timeout 300 python tools/devbench_cv.py 4736 > gpurun_out/r2a_cv_tma8.log 2>&1
PLSB_GEMM_CW=16 timeout 300 python tools/devbench_cv.py 4736 > gpurun_out/r2a_cv_tma16.log 2>&1
PLSB_GEMM_V1=1 timeout 300 python tools/devbench_cv.py 4736 > gpurun_out/r2a_cv_v1.log 2>&1
timeout 900 python bench.py --steps 5 --warmup 3 --no-cpu > gpurun_out/r2a_bench.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"dmma_tma|cv_fold_fused" --launch-skip 20 --launch-count 4 -o gpurun_out/r2a_cv -f python tools/devbench_cv.py 4736 > gpurun_out/r2a_ncu_cv.log 2>&1
ls -la gpurun_out | grep r2a
