#!/bin/bash
# round 2: full GPU test suite (no -x), then ncu of the CV kernels
set -x
mkdir -p gpurun_out
timeout 2400 python -m pytest tests -m gpu -q --durations=15 > gpurun_out/r2b_pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r2b_pytest.log
