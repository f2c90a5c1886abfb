set -x
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1 --nproc-per-node 8"
SLIM="--gpus 8 --steps 20 --no-e2e --no-fwd-only --no-fold"
SSI_B200_TIMELINE=1 $TR --master-port 29591 bench.py $SLIM > gpurun_out/r02_bench_c5_n8_final2.json 2> gpurun_out/r02_bench_c5_n8_final2.err
grep -a '^{' gpurun_out/r02_bench_c5_n8_final2.json | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('RESULT', d['ms_per_step'], d['value'], d['warmup'], d['per_rank_kernel_ms_per_step'], d['parity'] and {k:v for k,v in d['parity'].items() if k!='how'}, d['step_ms_rank0'])"
grep -a "timeline rank" gpurun_out/r02_bench_c5_n8_final2.err | tr '}' '\n' | grep -a rank
