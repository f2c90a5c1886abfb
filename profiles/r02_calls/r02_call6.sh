set -x
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1 --nproc-per-node 8"
SLIM="--gpus 8 --steps 20 --no-e2e --no-fwd-only --no-fold --no-parity"
i=0
for ex in peer nvls peer nvls; do
  i=$((i+1))
  SSI_B200_TIMELINE=1 $TR --master-port $((29530+i)) bench.py $SLIM --exchange $ex > gpurun_out/r02_diag_n8_${ex}_$i.json 2> gpurun_out/r02_diag_n8_${ex}_$i.err
  grep -a '^{' gpurun_out/r02_diag_n8_${ex}_$i.json | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$ex', d['ms_per_step'], d['per_rank_kernel_ms_per_step'], d['kernels']['exchange'])"
  grep -a "timeline" gpurun_out/r02_diag_n8_${ex}_$i.err
done
