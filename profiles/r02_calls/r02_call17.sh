set -x
CMD="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-e2e --no-fwd-only --no-gpu-reference --no-parity --no-fold"
for v in "base:" "store1:SSI_B200_STORE_HINT=1" "xlast:SSI_B200_K1_HINTS=20" "both:SSI_B200_STORE_HINT=1 SSI_B200_K1_HINTS=20"; do
  n=${v%%:*}; e=${v#*:}
  env $e ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none -k regex:gemm_kernel -s 60 -c 12 --csv --log-file gpurun_out/r02_k1_dram_$n.csv $CMD > /dev/null 2>&1
done
