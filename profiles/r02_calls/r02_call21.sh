set -x
CMD="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-e2e --no-fwd-only --no-gpu-reference --no-parity --no-fold"
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r02_launches_c5_final.csv $CMD > /dev/null 2>&1
ncu --set full --clock-control none --import-source on -k 'regex:gemm_kernel' -s 60 -c 6 -o gpurun_out/r02_prof_c5_gemm_final -f $CMD > gpurun_out/ncu_full_c5_final.log 2>&1
tail -2 gpurun_out/ncu_full_c5_final.log
