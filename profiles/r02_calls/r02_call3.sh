set -x
./tools/cluster_probe.bin | tee gpurun_out/r02_cluster_probe.json
nvidia-smi -L
python -m pytest tests -m gpu -x -q > gpurun_out/r02_pytest_gpu_2gpu.log 2>&1; tail -5 gpurun_out/r02_pytest_gpu_2gpu.log
python -m pytest tests/test_gpu_dp_peer.py tests/test_gpu_fold.py -m gpu -q -rA -s 2>&1 | tail -30 > gpurun_out/r02_pytest_dp_peer_2gpu.log; cat gpurun_out/r02_pytest_dp_peer_2gpu.log
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29513 tools/peer_probe.py > gpurun_out/r02_peer_probe_n2.txt 2>&1; tail -4 gpurun_out/r02_peer_probe_n2.txt
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 > gpurun_out/r02_bench_c5_n2.json 2> gpurun_out/r02_bench_c5_n2.err; tail -c 2500 gpurun_out/r02_bench_c5_n2.json; tail -5 gpurun_out/r02_bench_c5_n2.err
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --exchange nccl --no-e2e --no-fwd-only > gpurun_out/r02_bench_c5_n2_nccl.json 2> gpurun_out/r02_bench_c5_n2_nccl.err; tail -c 1200 gpurun_out/r02_bench_c5_n2_nccl.json
python bench.py --no-cpu-baseline --no-gpu-reference --no-e2e --no-fwd-only > gpurun_out/r02_bench_c5_n1_fold.json 2> gpurun_out/r02_bench_c5_n1_fold.err; tail -c 900 gpurun_out/r02_bench_c5_n1_fold.json
