set -x
L=speech-integration_b200/ssi_b200/libssi_b200.so
cp $L /tmp/base.so
for v in base a2 a3 m8 a3m8 a2m10 base; do
  if [ $v = base ]; then cp /tmp/base.so $L; else cp tools/_variants/lib_$v.so $L; fi
  for c in c3 c5; do
    python tools/bench_rows.py --config $c 2>/dev/null | python -c "
import sys,json
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); k=d['kernels']
print('$v $c', 'scatter', k['embedding_grad_scatter']['ms'], 'sort', k['embedding_grad_sort']['ms'], 'sorted', k['embedding_grad_scatter_sorted']['ms'])"
  done
done
cp /tmp/base.so $L
