#!/bin/bash
# ncu evidence for the final pass configuration of round 1 (one 512-thread CTA per SM, three-tile ring)
set -x
B="python bench.py --steps 2 --warmup 3 --no-cpu --no-cv --no-xtx --no-e2e"
$B > gpurun_out/r1e_plain.log 2>&1 || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r1e_launches_bench_c4.csv $B > gpurun_out/r1e_ncu_list.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:"fused_pass|xty_dmma|step_cluster|reduce_partials" --launch-skip 297 --launch-count 8 -o gpurun_out/r1e_fit -f $B > gpurun_out/r1e_ncu_fit.log 2>&1
ls -la gpurun_out | grep r1e
