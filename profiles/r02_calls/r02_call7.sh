set -x
python -m pytest tests/test_gpu_dp_peer.py tests/test_gpu_fold.py tests/test_abi.py -m gpu -q -x -s 2>&1 | tail -8
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1 --nproc-per-node 2"
$TR --master-port 29541 tools/sparse_probe.py > gpurun_out/r02_sparse_probe_n2_onephase.txt 2>&1; tail -4 gpurun_out/r02_sparse_probe_n2_onephase.txt
SSI_B200_SPARSE_TWO_PHASE=1 $TR --master-port 29542 tools/sparse_probe.py > gpurun_out/r02_sparse_probe_n2_twophase.txt 2>&1; tail -4 gpurun_out/r02_sparse_probe_n2_twophase.txt
