set -x
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1 --nproc-per-node 8"
SSI_B200_SPARSE_TWO_PHASE=0 $TR --master-port 29551 tools/sparse_probe.py > gpurun_out/r02_sparse_probe_n8_onephase.txt 2>&1; tail -9 gpurun_out/r02_sparse_probe_n8_onephase.txt | cut -c1-260
$TR --master-port 29552 tools/sparse_probe.py > gpurun_out/r02_sparse_probe_n8_twophase.txt 2>&1; tail -9 gpurun_out/r02_sparse_probe_n8_twophase.txt | cut -c1-260
SLIM="--gpus 8 --steps 20 --no-e2e --no-fwd-only --no-fold --no-parity"
i=0
for v in 0:peer -1:peer 0:peer -1:peer -1:nvls; do
  i=$((i+1)); tp=${v%%:*}; ex=${v##*:}
  SSI_B200_SPARSE_TWO_PHASE=$tp SSI_B200_TIMELINE=1 $TR --master-port $((29560+i)) bench.py $SLIM --exchange $ex > gpurun_out/r02_diag2_n8_${ex}_tp${tp}_$i.json 2> gpurun_out/r02_diag2_n8_${ex}_tp${tp}_$i.err
  grep -a '^{' gpurun_out/r02_diag2_n8_${ex}_tp${tp}_$i.json | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('RESULT $ex tp=$tp', d['ms_per_step'], d['per_rank_kernel_ms_per_step'], d['host_enqueue_ms_per_step'], d['parity'])"
  grep -a "timeline rank" gpurun_out/r02_diag2_n8_${ex}_tp${tp}_$i.err | tr '}' '\n' | grep -a "rank [03]"
done
