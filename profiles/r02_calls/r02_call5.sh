set -x
nvidia-smi -L | wc -l
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
$TR --nproc-per-node 8 --master-port 29521 bench.py --gpus 8 > gpurun_out/r02_bench_c5_n8.json 2> gpurun_out/r02_bench_c5_n8.err; tail -c 1500 gpurun_out/r02_bench_c5_n8.json; tail -3 gpurun_out/r02_bench_c5_n8.err
$TR --nproc-per-node 8 --master-port 29522 tools/peer_probe.py > gpurun_out/r02_peer_probe_n8.txt 2>&1; tail -3 gpurun_out/r02_peer_probe_n8.txt
SLIM="--no-e2e --no-fwd-only --no-fold --no-parity"
$TR --nproc-per-node 8 --master-port 29523 bench.py --gpus 8 --exchange nvls $SLIM > gpurun_out/r02_bench_c5_n8_nvls.json 2> gpurun_out/r02_bench_c5_n8_nvls.err; tail -c 700 gpurun_out/r02_bench_c5_n8_nvls.json
$TR --nproc-per-node 4 --master-port 29524 bench.py --gpus 4 > gpurun_out/r02_bench_c5_n4.json 2> gpurun_out/r02_bench_c5_n4.err; tail -c 700 gpurun_out/r02_bench_c5_n4.json
$TR --nproc-per-node 8 --master-port 29525 bench.py --gpus 8 --exchange nccl $SLIM > gpurun_out/r02_bench_c5_n8_nccl.json 2> gpurun_out/r02_bench_c5_n8_nccl.err; tail -c 700 gpurun_out/r02_bench_c5_n8_nccl.json
