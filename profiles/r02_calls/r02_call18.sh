set -x
python tools/ab.py --config c5 --steps 3 --rounds 4 base store_hint=1 store_hint=1,k1_group_m=64 store_hint=1,k1_group_m=32 store_hint=1,k1_hints=2 > gpurun_out/r02_ab_k1_l2hints2_c5.txt 2>&1; tail -6 gpurun_out/r02_ab_k1_l2hints2_c5.txt
CMD="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-e2e --no-fwd-only --no-gpu-reference --no-parity --no-fold"
for v in "s1g64:SSI_B200_STORE_HINT=1 SSI_B200_K1_GROUP_M=64" "s1g32:SSI_B200_STORE_HINT=1 SSI_B200_K1_GROUP_M=32" "s1w2:SSI_B200_STORE_HINT=1 SSI_B200_K1_HINTS=2"; do
  n=${v%%:*}; e=${v#*:}
  env $e ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none -k regex:gemm_kernel -s 60 -c 8 --csv --log-file gpurun_out/r02_k1_dram_$n.csv $CMD > /dev/null 2>&1
done
