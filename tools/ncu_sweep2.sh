M=dram__bytes_read.sum,dram__bytes_write.sum,lts__t_bytes.sum,gpu__time_duration.sum,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active
run() {
  tag=$1; shift
  env "$@" timeout 300 ncu --metrics $M --clock-control none -k regex:gemm_kernel -s 3 -c 3 --csv --log-file gpurun_out/sweep_$tag.csv python tools/one_step.py c3 2 > /dev/null 2>&1
  python tools/sweep_row.py "$tag"
}
for v in "$@"; do
  tag=$(echo "$v" | tr -c 'A-Za-z0-9\n' '_')
  run "$tag" $(echo "$v" | tr ',' ' ')
done
