# Round-1 profiling recipe, default workload c5 with 256x512 dW/dX tiles (run under gpurun, one GPU).
# Summaries: profiles/r01_c5_wide_*.md (first capture), r01_c5_final_*.md (after tail splitting + 4.5 GiB block), profiles/traffic.json
set -x
CMD="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-e2e"
$CMD > gpurun_out/plain_c5.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 420 --csv --log-file gpurun_out/launches_c5_final.csv $CMD > gpurun_out/ncu_launches_c5.log 2>&1
$CMD > gpurun_out/plain_c5b.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k 'regex:gemm_kernel|softmax_grad' -s 32 -c 8 -o gpurun_out/prof_c5_final -f $CMD > gpurun_out/ncu_full_c5.log 2>&1
tail -3 gpurun_out/ncu_full_c5.log
