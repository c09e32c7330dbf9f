#!/bin/bash
# GPU session 60 (1 GPU): one-piece-per-thread RoPE + cache write, RowDivider quantisers, vectorised reduce_quant;
# silu_mul_quant split sweep; e2e host-time breakdown.
mkdir -p gpurun_out; S=gpurun_out/r60_summary.txt; : > $S
timeout 900 python -m pytest tests/test_gpu_layer_glue.py tests/test_gpu_quant.py tests/test_gpu_attention.py tests/test_gpu_cache.py tests/test_gpu_decode_step.py -m gpu -q --timeout 300 -x > gpurun_out/r60_tests.log 2>&1; echo "tests exit $?" >> $S
tail -4 gpurun_out/r60_tests.log >> $S
grep -E "^FAILED|^ERROR" gpurun_out/r60_tests.log | head -20 >> $S
timeout 300 python benchmarks/bench_layer_glue.py --tokens 256 > gpurun_out/r60_glue_256.jsonl 2> gpurun_out/r60_glue_256.err; echo "glue 256 exit $?" >> $S
timeout 300 python benchmarks/bench_layer_glue.py --tokens 16 > gpurun_out/r60_glue_16.jsonl 2> gpurun_out/r60_glue_16.err; echo "glue 16 exit $?" >> $S
for sp in 1 2 4 8; do
  echo "silu split $sp" >> $S
  VB200_SILU_SPLIT=$sp timeout 300 python benchmarks/bench_layer_glue.py --tokens 256 2>/dev/null | grep silu_and_mul_dynamic >> $S
done
timeout 600 python bench.py --no-cpu-baseline > gpurun_out/r60_bench_default.log 2>&1; echo "bench default exit $?" >> $S
cat gpurun_out/r60_glue_256.jsonl gpurun_out/r60_glue_16.jsonl >> $S
grep -h '^{' gpurun_out/r60_bench_default.log | python -c "
import sys, json
d = json.loads(sys.stdin.read())
print({k: d[k] for k in ('value', 'ms_per_step', 'e2e', 'gpu_launches_per_step', 'clocks')})
print(d['roofline']['avg_launch_us'], d['roofline']['frac'], d['roofline_gemm']['sum_us_per_layer'])" >> $S
cat $S
