set -x
python -m pytest tests -m gpu -x -q > gpurun_out/r02_pytest_gpu_1gpu.log 2>&1; tail -4 gpurun_out/r02_pytest_gpu_1gpu.log
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
