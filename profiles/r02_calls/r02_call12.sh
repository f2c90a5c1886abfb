set -x
python -m pytest tests -m gpu -x -q > gpurun_out/r02_pytest_gpu_1gpu.log 2>&1; tail -3 gpurun_out/r02_pytest_gpu_1gpu.log
python bench.py --no-cpu-baseline --no-gpu-reference --no-e2e --no-fwd-only --no-fold > gpurun_out/r02_bench_c5_n1_scan.json 2>/dev/null
python -c "
import json
d=json.loads([l for l in open('gpurun_out/r02_bench_c5_n1_scan.json') if l.startswith('{')][-1])
print(d['ms_per_step'], {k:round(v['ms_per_step'],2) for k,v in d['kernels'].items()})"
