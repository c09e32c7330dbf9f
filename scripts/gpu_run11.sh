#!/bin/bash
# GPU session 11: bulk-copy (TMA engine) paged attention: parity, ring-depth sweep, bench; GEMM trace.
mkdir -p gpurun_out; S=gpurun_out/r11_summary.txt; : > $S
timeout 600 python -m pytest tests/test_gpu_attention.py tests/test_gpu_decode_step.py -m gpu -q --timeout 120 > gpurun_out/r11_pa_tests.log 2>&1; echo "pa tests exit $?" >> $S
tail -5 gpurun_out/r11_pa_tests.log >> $S
for st in 0 4 8 12; do
  VB200_PA_STAGES=$st timeout 300 python benchmarks/bench_paged_attention.py --no-ref --iters 30 > gpurun_out/r11_pa_st$st.log 2>&1; echo "stages $st exit $?" >> $S
done
VB200_PA_TMA=0 timeout 300 python benchmarks/bench_paged_attention.py --no-ref --iters 30 > gpurun_out/r11_pa_old.log 2>&1
timeout 300 python benchmarks/bench_paged_attention.py --iters 20 > gpurun_out/r11_pa_ref.log 2>&1; echo "pa with ref exit $?" >> $S
timeout 900 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/r11_bench_cfg3.log 2>&1; echo "bench exit $?" >> $S
timeout 300 python benchmarks/trace_scaled_mm.py > gpurun_out/r11_mm_trace.log 2>&1; echo "trace exit $?" >> $S
cat $S
