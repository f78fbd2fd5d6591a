#!/bin/bash
# ncu evidence for round 1d (run on the GPU box through gpurun; reports land in gpurun_out/, summaries in profiles/)
set -x
B="python bench.py --steps 2 --warmup 3 --no-cpu --no-cv --no-xtx --e2e-steps 1"
$B > gpurun_out/r1d_plain.log 2>&1 || exit 1
# launch list of the device-resident fits AND the end-to-end path (upload with X'X / X'Y under the transfer, fit from X'X, T = X R)
ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file gpurun_out/r1d_launches_bench_c4.csv $B > gpurun_out/r1d_ncu_list.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:"fused_pass|xty_dmma" --launch-skip 165 --launch-count 2 -o gpurun_out/r1d_fit -f $B > gpurun_out/r1d_ncu_fit.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:"scores_dmma|symv_kernel" --launch-skip 30 --launch-count 3 -o gpurun_out/r1d_e2e -f $B > gpurun_out/r1d_ncu_e2e.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:"dmma_atb" --launch-skip 100 --launch-count 1 -o gpurun_out/r1d_syrk -f $B > gpurun_out/r1d_ncu_syrk.log 2>&1
C="python bench.py --steps 1 --warmup 3 --no-e2e --no-cpu --no-xtx"
ncu --set full --clock-control none --import-source on -k regex:"cv_fold_fused|dmma_atb" --launch-skip 150 --launch-count 4 -o gpurun_out/r1d_cv -f $C > gpurun_out/r1d_ncu_cv.log 2>&1
ls -la gpurun_out
