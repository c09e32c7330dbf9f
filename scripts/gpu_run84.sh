#!/bin/bash
# GPU session 84 (1 GPU): ncu counters behind the grouped-GEMM changes at w13 / batch 512 (ragged): DRAM bytes and L2 -> SM
# bytes per launch with the old launch order, the new one, and the new one with two weight tiles per CTA.
mkdir -p gpurun_out; S=gpurun_out/r84_summary.txt; : > $S
M="gpu__time_duration.sum,dram__bytes_read.sum,lts__t_sectors_srcunit_tex_op_read.sum,l1tex__m_xbar2l1tex_read_bytes.sum,sm__cycles_active.avg,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed,launch__grid_size,launch__occupancy_limit_shared_mem"
cap() {  # name, env...
  local name=$1; shift
  env "$@" timeout 300 ncu --metrics $M --clock-control none -k regex:moe_grouped_gemm_kernel --launch-skip 4 -c 2 --csv --log-file gpurun_out/r84_$name.csv python benchmarks/bench_moe.py --tps 1 --batches 512 > gpurun_out/r84_$name.log 2>&1
  echo "## $name (exit $?)" >> $S
  python - gpurun_out/r84_$name.csv <<'PY' >> gpurun_out/r84_summary.txt
import csv, sys
rows = [r for r in csv.reader(l for l in open(sys.argv[1]) if l.startswith('"'))]
hdr = rows[0]; i_id, i_k, i_m, i_u, i_v = (hdr.index(x) for x in ("ID", "Kernel Name", "Metric Name", "Metric Unit", "Metric Value"))
for r in rows[1:]:
    if r[i_id] == "0":
        print("  %-70s %16s %s" % (r[i_m], r[i_v], r[i_u]))
print("  kernel:", rows[1][i_k][:90])
PY
}
cap order0_nt1 VB200_MOE_ORDER=0 VB200_MOE_NT=1
cap order1_nt1 VB200_MOE_NT=1
cap order1_nt2 VB200_MOE_NT=2
cat $S
