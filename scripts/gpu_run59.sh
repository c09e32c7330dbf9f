#!/bin/bash
# GPU session 59 (1 GPU): greedy_argmax kernel + packed / double-buffered e2e inputs; in-graph cost of the small
# per-token kernels (benchmarks/bench_layer_glue.py) at decode batch 256 and 16.
mkdir -p gpurun_out; S=gpurun_out/r59_summary.txt; : > $S
timeout 600 python -m pytest tests/test_gpu_decode_loop.py tests/test_gpu_decode_step.py -m gpu -q --timeout 300 > gpurun_out/r59_tests.log 2>&1; echo "tests exit $?" >> $S
tail -4 gpurun_out/r59_tests.log >> $S
grep -E "^FAILED|^ERROR" gpurun_out/r59_tests.log | head -20 >> $S
timeout 300 python benchmarks/bench_layer_glue.py --tokens 256 > gpurun_out/r59_glue_256.jsonl 2> gpurun_out/r59_glue_256.err; echo "glue 256 exit $?" >> $S
timeout 300 python benchmarks/bench_layer_glue.py --tokens 16 > gpurun_out/r59_glue_16.jsonl 2> gpurun_out/r59_glue_16.err; echo "glue 16 exit $?" >> $S
timeout 600 python bench.py --no-cpu-baseline > gpurun_out/r59_bench_default.log 2>&1; echo "bench default exit $?" >> $S
cat gpurun_out/r59_glue_256.jsonl gpurun_out/r59_glue_16.jsonl >> $S
tail -3 gpurun_out/r59_glue_256.err >> $S
grep -h '^{' gpurun_out/r59_bench_default.log | cut -c1-1200 >> $S
cat $S
