#!/bin/bash
# last check of the round on the committed tree: GPU suite, smoke, short bench
set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x --durations=5 > gpurun_out/final_pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/final_pytest.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/final_smoke.log 2>&1
echo "smoke rc=$?" >> gpurun_out/final_smoke.log
timeout 1200 python bench.py --steps 10 --warmup 3 --no-cpu > gpurun_out/final_bench.log 2>&1
echo "bench rc=$?" >> gpurun_out/final_bench.log
