#!/bin/bash
# GPU session 53 (1 GPU): scaled_mm with running (stage, phase) counters in the single-thread producer / MMA loops.
mkdir -p gpurun_out; S=gpurun_out/r53_summary.txt; : > $S
timeout 900 python -m pytest tests/test_gpu_scaled_mm.py -m gpu -q --timeout 300 -x > gpurun_out/r53_tests.log 2>&1; echo "mm tests exit $?" >> $S
tail -3 gpurun_out/r53_tests.log >> $S
timeout 300 python benchmarks/trace_scaled_mm.py > gpurun_out/r53_mm_trace.log 2>&1; echo "trace exit $?" >> $S
timeout 600 python benchmarks/bench_scaled_mm.py > gpurun_out/r53_bench_mm.log 2>&1; echo "bench_mm exit $?" >> $S
cat $S
timeout 600 python bench.py --no-cpu-baseline > gpurun_out/r53_bench_default.log 2>&1; echo "bench exit $?"
grep -h '^{' gpurun_out/r53_bench_default.log | cut -c1-300
