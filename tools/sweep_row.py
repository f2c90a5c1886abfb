import csv, sys
tag = sys.argv[1]
rows = list(csv.reader(open(f"gpurun_out/sweep_{tag}.csv", errors="replace")))
h = next(i for i, r in enumerate(rows) if r and r[0] == "ID")
hdr = rows[h]; ki, mi, vi = hdr.index("Kernel Name"), hdr.index("Metric Name"), hdr.index("Metric Value")
out = {}
for r in rows[h + 1:]:
    if len(r) <= vi: continue
    out.setdefault((r[0], r[ki][17:28]), {})[r[mi]] = float(r[vi].replace(",", ""))
for (i, k), m in out.items():
    print(f"{tag:28s} {k} rd {m['dram__bytes_read.sum']/1e9:6.2f} GB wr {m['dram__bytes_write.sum']/1e9:5.2f} GB l2 {m['lts__t_bytes.sum']/1e9:6.1f} GB "
          f"t {m['gpu__time_duration.sum']/1e6:5.2f} ms tensor {m['sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active']:5.1f}%", flush=True)
