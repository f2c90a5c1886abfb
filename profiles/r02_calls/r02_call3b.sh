set -x
./tools/cluster_probe.bin | tee gpurun_out/r02_cluster_probe.json
python -m pytest tests/test_gpu_fold.py tests/test_gpu_dp_peer.py tests/test_abi.py -m gpu -q -x 2>&1 | tail -15
python -m pytest tests -m gpu -x -q 2>&1 | tail -3
python bench.py --no-cpu-baseline --no-gpu-reference --no-e2e --no-fwd-only > gpurun_out/r02_bench_c5_n1_fold.json 2> gpurun_out/r02_bench_c5_n1_fold.err; tail -c 900 gpurun_out/r02_bench_c5_n1_fold.json; tail -3 gpurun_out/r02_bench_c5_n1_fold.err
