set -x
python -m pytest tests -m gpu -x -q > gpurun_out/r02_pytest_gpu_1gpu.log 2>&1; tail -3 gpurun_out/r02_pytest_gpu_1gpu.log
python tools/bench_rows.py --config c3 > gpurun_out/r02_rows_c3.json 2>/dev/null; python tools/bench_rows.py --config c5 > gpurun_out/r02_rows_c5.json 2>/dev/null
python -c "
import json
for c in ('c3','c5'):
    d=json.loads(open('gpurun_out/r02_rows_%s.json'%c).read().strip().splitlines()[-1]); print(c, {k:(v['ms'],v['frac_of_hbm_peak']) for k,v in d['kernels'].items()}, d['eval_forward_only']['ms'])"
