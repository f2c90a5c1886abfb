# DRAM / L2 traffic of the three contractions for tuning variants (ncu metrics are deterministic, timings are not)
M=dram__bytes_read.sum,dram__bytes_write.sum,lts__t_bytes.sum,gpu__time_duration.sum,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active
run() {
  tag=$1; shift
  env "$@" ncu --metrics $M --clock-control none -k regex:gemm_kernel -s 3 -c 3 --csv --log-file gpurun_out/sweep_$tag.csv python tools/one_step.py c3 2 > /dev/null 2>&1
  python - "$tag" <<'P'
import csv, sys
tag = sys.argv[1]
rows = list(csv.reader(open(f"gpurun_out/sweep_{tag}.csv", errors="replace")))
h = next(i for i, r in enumerate(rows) if r and r[0] == "ID")
hdr = rows[h]; ki, mi, vi, ui = hdr.index("Kernel Name"), hdr.index("Metric Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
out = {}
for r in rows[h + 1:]:
    if len(r) <= vi: continue
    out.setdefault((r[0], r[ki][:28]), {})[r[mi]] = (r[vi], r[ui])
for (i, k), m in out.items():
    print(tag, k, " ".join(f"{n.split('__')[1][:22]}={v}{u}" for n, (v, u) in m.items()))
P
}
run base A=1
run cg1 SSI_B200_CTA_GROUP=1
run dw12 SSI_B200_DW_HINTS=12
run dw21 SSI_B200_DW_HINTS=21
run dx12 SSI_B200_DX_HINTS=12
run dx21 SSI_B200_DX_HINTS=21
run dx10 SSI_B200_DX_HINTS=10
run k121 SSI_B200_K1_HINTS=21
run k121s SSI_B200_K1_HINTS=21 SSI_B200_STORE_HINT=1
run gdx4 SSI_B200_GROUP_M_DX=4
run gdx8 SSI_B200_GROUP_M_DX=8
run gdx38 SSI_B200_GROUP_M_DX=38
run gdx74 SSI_B200_GROUP_M_DX=74
run gdw4 SSI_B200_GROUP_M_DW=4
run gdw8 SSI_B200_GROUP_M_DW=8
run gdw38 SSI_B200_GROUP_M_DW=38
run gdw74 SSI_B200_GROUP_M_DW=74
