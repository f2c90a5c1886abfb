set -x
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1 --nproc-per-node 8"
$TR --master-port 29571 bench.py --gpus 8 > gpurun_out/r02_bench_c5_n8_final.json 2> gpurun_out/r02_bench_c5_n8_final.err; tail -c 600 gpurun_out/r02_bench_c5_n8_final.json; tail -2 gpurun_out/r02_bench_c5_n8_final.err
SLIM="--gpus 8 --steps 20 --no-e2e --no-fwd-only --no-fold --no-parity"
SSI_B200_TIMELINE=1 $TR --master-port 29572 bench.py $SLIM > gpurun_out/r02_diag3_n8.json 2> gpurun_out/r02_diag3_n8.err
grep -a '^{' gpurun_out/r02_diag3_n8.json | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('RESULT', d['ms_per_step'], d['warmup'], d['per_rank_kernel_ms_per_step'], d['host_enqueue_ms_per_step'], d['step_ms_rank0'])"
grep -a "timeline rank" gpurun_out/r02_diag3_n8.err | tr '}' '\n' | grep -a "rank"
