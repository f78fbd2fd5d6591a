#!/bin/bash
# round 2, session 4: full GPU suite, small-call trace, full bench line (with the CPU legs)
set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x --durations=8 > gpurun_out/r2n_pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r2n_pytest.log
timeout 300 python tools/trace_c3.py > gpurun_out/r2n_trace_c3.log 2>&1
timeout 1800 python bench.py --steps 10 --warmup 3 > gpurun_out/r2n_bench.log 2>&1
