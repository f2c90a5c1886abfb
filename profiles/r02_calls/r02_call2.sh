set -x
python tools/ab.py --config c5 --steps 3 --rounds 3 base k1_group_m=64 k1_group_m=32 > gpurun_out/r02_ab_k1_group_c5.txt 2>&1; tail -12 gpurun_out/r02_ab_k1_group_c5.txt
python bench.py --config c3 --mode recompute --no-cpu-baseline --no-gpu-reference > gpurun_out/r02_bench_c3_recompute.json 2> gpurun_out/r02_bench_c3_recompute.err; tail -c 1500 gpurun_out/r02_bench_c3_recompute.json
python bench.py --config c5 --mode recompute --no-cpu-baseline --no-gpu-reference > gpurun_out/r02_bench_c5_recompute.json 2> gpurun_out/r02_bench_c5_recompute.err; tail -c 1500 gpurun_out/r02_bench_c5_recompute.json
python -m pytest tests -m gpu -x -q -k "recompute or bwd or parity" 2>&1 | tail -3
