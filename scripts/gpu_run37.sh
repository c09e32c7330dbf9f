#!/bin/bash
# GPU session 37 (1 GPU): L2 / HBM -> shared-memory feed probe (unicast vs cluster multicast), cuBLASLt kernel
# configurations at decode M (ncu launch list), planner fix check.
mkdir -p gpurun_out; S=gpurun_out/r37_summary.txt; : > $S
P=benchmarks/bin/l2_feed_probe; O=gpurun_out/r37_l2_feed_probe.jsonl; : > $O
for cfg in "148 1 0 0" "96 1 0 0" "74 1 0 0" "37 1 0 0" "8 1 0 0" "148 2 0 0" "148 2 1 0" "132 4 0 0" "132 4 1 0" "128 8 0 0" "128 8 1 0" \
           "148 1 0 1" "96 1 0 1" "148 2 1 1" "132 4 1 1" "148 1 0 0 32 6" "148 2 1 0 32 6" "132 4 1 0 32 6" "148 1 0 1 32 6" "148 1 0 0 8 16" "148 1 0 0 4 16"; do
  timeout 60 $P $cfg >> $O 2>&1 || echo "probe $cfg failed $?" >> $S
done
cat $O >> $S
timeout 600 ncu --metrics gpu__time_duration.sum,launch__grid_size,launch__block_size,launch__cluster_size,launch__shared_mem_per_block_dynamic,launch__registers_per_thread --clock-control none --csv --log-file gpurun_out/r37_cublaslt_kernels.csv python benchmarks/probe_cublaslt_kernels.py > gpurun_out/r37_probe_cublaslt.log 2>&1; echo "cublaslt probe exit $?" >> $S
timeout 600 python benchmarks/bench_scaled_mm.py --shapes gate_up --M 128,256 > gpurun_out/r37_bench_mm_gate_up.log 2>&1; echo "bench_mm exit $?" >> $S
grep '^{' gpurun_out/r37_bench_mm_gate_up.log | cut -c1-200 >> $S
cat $S
