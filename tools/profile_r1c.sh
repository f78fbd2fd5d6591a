set -x
B="python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu --no-cv"
$B > gpurun_out/r1c_plain.log 2>&1 || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -c 500 --csv --log-file gpurun_out/r1c_launches_bench_c4.csv $B > gpurun_out/r1c_ncu_list.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:"fused_pass|xty_dmma|step_cluster|reduce_partials" --launch-skip 297 --launch-count 8 -o gpurun_out/r1c_fit -f $B > gpurun_out/r1c_ncu_fit.log 2>&1
C="python bench.py --steps 1 --warmup 3 --no-e2e --no-cpu"
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"cv_fold|dmma_atb|gemm_reduce" -c 300 --csv --log-file gpurun_out/r1c_launches_cv_c5.csv $C > gpurun_out/r1c_ncu_cvlist.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:"cv_fold|dmma_atb" --launch-skip 40 --launch-count 4 -o gpurun_out/r1c_cv -f $C > gpurun_out/r1c_ncu_cv.log 2>&1
ls -la gpurun_out
