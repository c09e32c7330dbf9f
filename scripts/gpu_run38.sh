#!/bin/bash
# GPU session 38 (1 GPU): three TMA producer warps in scaled_mm: parity, sweep vs torch._scaled_mm and the reference's sm100 CUTLASS kernel,
# per-phase trace; feed probe with 1/2/4 issuing threads.
mkdir -p gpurun_out; S=gpurun_out/r38_summary.txt; : > $S
P=benchmarks/bin/l2_feed_probe; O=gpurun_out/r38_l2_feed_probe.jsonl; : > $O
for cfg in "148 1 0 0 16 8 1" "148 1 0 0 16 8 2" "148 1 0 0 16 8 4" "148 1 0 0 8 16 1" "148 1 0 0 8 16 2" "148 1 0 0 8 16 4" "148 1 0 0 4 16 4" "148 1 0 0 4 16 8" "148 1 0 1 16 8 1" "148 1 0 1 16 8 4" "96 1 0 1 16 8 4"; do
  timeout 60 $P $cfg >> $O 2>&1 || echo "probe $cfg failed $?" >> $S
done
cat $O >> $S
timeout 900 python -m pytest tests/test_gpu_scaled_mm.py tests/test_gpu_decode_step.py -m gpu -q --timeout 300 -x > gpurun_out/r38_tests.log 2>&1; echo "mm+decode tests exit $?" >> $S
tail -4 gpurun_out/r38_tests.log >> $S
timeout 600 python benchmarks/bench_scaled_mm.py > gpurun_out/r38_bench_mm.log 2>&1; echo "bench_mm exit $?" >> $S
timeout 300 python benchmarks/trace_scaled_mm.py > gpurun_out/r38_mm_trace.log 2>&1; echo "trace exit $?" >> $S
timeout 600 python bench.py --no-cpu-baseline > gpurun_out/r38_bench_default.log 2>&1; echo "bench default exit $?" >> $S
grep -h '^{' gpurun_out/r38_bench_default.log | cut -c1-300 >> $S
cat $S
