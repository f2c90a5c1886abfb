set -x
python tools/ab.py --config c5 --steps 3 --rounds 3 base store_hint=1 k1_hints=20 store_hint=1,k1_hints=20 k1_hints=22 > gpurun_out/r02_ab_k1_l2hints_c5.txt 2>&1; tail -6 gpurun_out/r02_ab_k1_l2hints_c5.txt
CMD="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-e2e --no-fwd-only --no-gpu-reference --no-parity --no-fold"
for v in "base:" "store1:SSI_B200_STORE_HINT=1" "xlast:SSI_B200_K1_HINTS=20" "both:SSI_B200_STORE_HINT=1 SSI_B200_K1_HINTS=20"; do
  n=${v%%:*}; e=${v#*:}
  env $e ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none --kernel-name-base demangled -k 'regex:gemm_kernel<5' -s 12 -c 3 --csv --log-file gpurun_out/r02_k1_dram_$n.csv $CMD > /dev/null 2>&1
  python - <<PY
import csv
rows=list(csv.reader(open('gpurun_out/r02_k1_dram_$n.csv', errors='replace')))
h=next(i for i,r in enumerate(rows) if r and r[0]=='ID'); hdr=rows[h]
mi,vi,ui=hdr.index('Metric Name'),hdr.index('Metric Value'),hdr.index('Metric Unit')
agg={}
for r in rows[h+1:]:
    if len(r)>vi: agg.setdefault(r[mi],[]).append((r[vi],r[ui]))
print('$n', agg)
PY
done
