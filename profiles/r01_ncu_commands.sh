# Round-1 profiling recipe (run under gpurun, one GPU). Summaries: profiles/r01_*_launches.md, r01_*_ncu_full.md
set -x
CMD="python bench.py --steps 1 --warmup 3 --config c3 --no-cpu-baseline --no-e2e"
$CMD > gpurun_out/plain_c3.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches_c3.csv $CMD > gpurun_out/ncu_launches.log 2>&1
$CMD > gpurun_out/plain_c3b.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k 'regex:gemm_kernel|softmax_grad' -s 24 -c 8 -o gpurun_out/prof_c3 $CMD > gpurun_out/ncu_full.log 2>&1
tail -3 gpurun_out/ncu_full.log
