set -x
python -m pytest tests -m gpu -x -q > gpurun_out/r02_pytest_gpu_1gpu.log 2>&1; tail -3 gpurun_out/r02_pytest_gpu_1gpu.log
python bench.py > gpurun_out/r02_bench_c5_n1_final2.json 2> gpurun_out/r02_bench_c5_n1_final2.err; tail -c 300 gpurun_out/r02_bench_c5_n1_final2.json
CMD="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-e2e --no-fwd-only --no-gpu-reference --no-parity --no-fold"
ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum,l1tex__m_xbar2l1tex_read_bytes.sum,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active --clock-control none -k regex:gemm_kernel -s 60 -c 6 --csv --log-file gpurun_out/r02_gemm_metrics_final.csv $CMD > /dev/null 2>&1
