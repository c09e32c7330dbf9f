#!/bin/bash
# GPU session 64 (1 GPU): rms_norm sized to the row + weight prefetch, silu split 8; glue bench + config 3 bench.
mkdir -p gpurun_out; S=gpurun_out/r64_summary.txt; : > $S
timeout 900 python -m pytest tests/test_gpu_layer_glue.py tests/test_gpu_quant.py tests/test_gpu_decode_step.py -m gpu -q --timeout 300 -x > gpurun_out/r64_tests.log 2>&1; echo "tests exit $?" >> $S
tail -4 gpurun_out/r64_tests.log >> $S
timeout 300 python benchmarks/bench_layer_glue.py --tokens 256 > gpurun_out/r64_glue_256.jsonl 2> gpurun_out/r64_glue_256.err; echo "glue 256 exit $?" >> $S
timeout 300 python benchmarks/bench_layer_glue.py --tokens 16 > gpurun_out/r64_glue_16.jsonl 2> gpurun_out/r64_glue_16.err; echo "glue 16 exit $?" >> $S
head -7 gpurun_out/r64_glue_256.jsonl >> $S; head -7 gpurun_out/r64_glue_16.jsonl >> $S
timeout 600 python bench.py --no-cpu-baseline > gpurun_out/r64_bench_default.log 2>&1; echo "bench default exit $?" >> $S
grep -h '^{' gpurun_out/r64_bench_default.log | python -c "
import sys, json
d = json.loads(sys.stdin.read())
print({k: d[k] for k in ('value', 'ms_per_step', 'e2e', 'gpu_launches_per_step', 'clocks')})
print(d['roofline']['avg_launch_us'], d['roofline']['frac'], d['roofline_gemm']['sum_us_per_layer'])" >> $S
cat $S
