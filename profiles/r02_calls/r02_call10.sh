set -x
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
$TR --nproc-per-node 4 --master-port 29581 bench.py --gpus 4 > gpurun_out/r02_bench_c5_n4_final.json 2> gpurun_out/r02_bench_c5_n4_final.err; tail -c 400 gpurun_out/r02_bench_c5_n4_final.json
$TR --nproc-per-node 2 --master-port 29582 bench.py --gpus 2 > gpurun_out/r02_bench_c5_n2_final.json 2> gpurun_out/r02_bench_c5_n2_final.err; tail -c 400 gpurun_out/r02_bench_c5_n2_final.json
python bench.py > gpurun_out/r02_bench_c5_n1_final.json 2> gpurun_out/r02_bench_c5_n1_final.err; tail -c 400 gpurun_out/r02_bench_c5_n1_final.json
