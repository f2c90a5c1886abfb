# Round-2 profiling recipe (run under gpurun, one GPU). Summaries: profiles/r02_c5_*.md, r02_k3_ncu.md, r02_sanitizer_refused.log (the pool refuses sanitizer runs; the two commands were removed)
set -x
python -m pytest tests -m gpu -x -q > gpurun_out/r02_pytest_gpu.log 2>&1; tail -3 gpurun_out/r02_pytest_gpu.log
python bench.py > gpurun_out/r02_bench_c5_n1.json 2> gpurun_out/r02_bench_c5_n1.err; tail -c 600 gpurun_out/r02_bench_c5_n1.json
CMD="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-e2e --no-fwd-only --no-gpu-reference --no-parity --no-fold"
$CMD > gpurun_out/plain_c5.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r02_launches_c5.csv $CMD > gpurun_out/ncu_launches_c5.log 2>&1
ncu --set full --clock-control none --import-source on -k 'regex:gemm_kernel' -s 40 -c 6 -o gpurun_out/r02_prof_c5_gemm -f $CMD > gpurun_out/ncu_full_c5.log 2>&1
tail -3 gpurun_out/ncu_full_c5.log
python tools/bench_rows.py --config c3 > gpurun_out/r02_rows_c3.json 2>gpurun_out/r02_rows_c3.err &&
ncu --set full --clock-control none --import-source on -k 'regex:sort|scatter' -c 12 -o gpurun_out/r02_prof_k3 -f python tools/bench_rows.py --config c3 --iters 1 > gpurun_out/ncu_full_k3.log 2>&1
tail -3 gpurun_out/ncu_full_k3.log
