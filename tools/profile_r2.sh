#!/bin/bash
# round 2 profiles: ncu launch lists (shares) and --set full captures of the kernels this round added or changed.
# Every profiled command first runs once WITHOUT ncu (exit code checked); numbers printed under ncu are never bench values.
set -x
mkdir -p gpurun_out
B="python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu --no-cv"
$B > gpurun_out/r2p_bench_plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 900 --csv --log-file gpurun_out/r2p_launches_bench_c4.csv $B > gpurun_out/r2p_ncu_bench.log 2>&1
C="python tools/devbench_cv2.py 7000 1"
$C > gpurun_out/r2p_cv_plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 1200 --csv --log-file gpurun_out/r2p_launches_cv_c5.csv $C > gpurun_out/r2p_ncu_cvl.log 2>&1
timeout 900 ncu --set full --clock-control none -k regex:"cv_gemm_kernel|cv_fold_fused" --launch-skip 330 --launch-count 4 -o gpurun_out/r2p_cv -f $C > gpurun_out/r2p_ncu_cv.log 2>&1
timeout 900 ncu --set full --clock-control none -k regex:"fused_pass_kernel|xty_dmma_kernel|dmma_tma_kernel|step_cluster_kernel|reduce_partials_kernel" --launch-skip 150 --launch-count 6 -o gpurun_out/r2p_fit -f $B > gpurun_out/r2p_ncu_fit.log 2>&1
timeout 900 ncu --set full --clock-control none -k regex:"dmma_tma_kernel|gemm_reduce" --launch-count 3 -o gpurun_out/r2p_xtx -f $B > gpurun_out/r2p_ncu_xtx.log 2>&1
S="python tools/devbench_cv.py 256"
$S > gpurun_out/r2p_small_plain.log 2>&1 && \
timeout 600 ncu --set full --clock-control none -k regex:"cv_warp_kernel|cv_small_kernel" --launch-count 6 -o gpurun_out/r2p_small -f $S > gpurun_out/r2p_ncu_small.log 2>&1
ls -la gpurun_out | grep r2p
# the reports are large: keep their raw metric pages as CSV (what tools/ncu_summary.py reads) and drop the .ncu-rep files
for r in gpurun_out/r2p_*.ncu-rep; do ncu -i $r --page raw --csv > ${r%.ncu-rep}_raw.csv 2>/dev/null; rm -f $r; done
du -sh gpurun_out
