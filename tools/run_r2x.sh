#!/bin/bash
# final validation of the round: GPU suite, smoke, full bench line (with CPU legs), reference arm (full-size fit)
set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x --durations=5 > gpurun_out/r2x_pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r2x_pytest.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2x_smoke.log 2>&1
echo "smoke rc=$?" >> gpurun_out/r2x_smoke.log
timeout 1800 python bench.py > gpurun_out/r2x_bench.log 2>&1
echo "bench rc=$?" >> gpurun_out/r2x_bench.log
timeout 1500 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r2x_bench_reference.log 2>&1
echo "reference rc=$?" >> gpurun_out/r2x_bench_reference.log
