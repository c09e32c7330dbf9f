#!/bin/bash
# GPU session 13: as 12 with the shared-memory alignment fix (GEMM a_scale tile, MoE token ids).
mkdir -p gpurun_out; S=gpurun_out/r13_summary.txt; : > $S
timeout 1500 python -m pytest tests -m gpu -q --timeout 120 > gpurun_out/r13_tests.log 2>&1; echo "tests exit $?" >> $S
tail -5 gpurun_out/r13_tests.log >> $S
timeout 300 python benchmarks/bench_paged_attention.py --no-ref --iters 30 > gpurun_out/r13_pa.log 2>&1; echo "pa exit $?" >> $S
for st in 8; do
  VB200_PA_STAGES=$st timeout 300 python benchmarks/bench_paged_attention.py --no-ref --iters 30 > gpurun_out/r13_pa_st$st.log 2>&1; echo "stages $st exit $?" >> $S
done
timeout 300 python benchmarks/bench_scaled_mm.py --M 16,64,128,256,1024,4096 > gpurun_out/r13_bench_mm.log 2>&1; echo "bench_mm exit $?" >> $S
timeout 300 python benchmarks/trace_scaled_mm.py > gpurun_out/r13_mm_trace.log 2>&1; echo "trace exit $?" >> $S
timeout 900 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/r13_bench_cfg3.log 2>&1; echo "bench exit $?" >> $S
cat $S
