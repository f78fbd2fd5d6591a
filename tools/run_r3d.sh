#!/bin/bash
# final state of the round: GPU suite, smoke, bench line, ncu profiles of the final code
set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x --durations=5 > gpurun_out/r3d_pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r3d_pytest.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r3d_smoke.log 2>&1
echo "smoke rc=$?" >> gpurun_out/r3d_smoke.log
timeout 1800 python bench.py > gpurun_out/r3d_bench.log 2>&1
echo "bench rc=$?" >> gpurun_out/r3d_bench.log
bash tools/profile_r2.sh
