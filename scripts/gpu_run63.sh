#!/bin/bash
# GPU session 63 (1 GPU): silu_mul_quant with registers sized to the slice (one wave) + paired rounding.
mkdir -p gpurun_out; S=gpurun_out/r63_summary.txt; : > $S
timeout 900 python -m pytest tests/test_gpu_layer_glue.py tests/test_gpu_quant.py tests/test_gpu_decode_step.py -m gpu -q --timeout 300 -x > gpurun_out/r63_tests.log 2>&1; echo "tests exit $?" >> $S
tail -4 gpurun_out/r63_tests.log >> $S
for sp in 1 2 4 8; do
  echo "silu split $sp" >> $S
  VB200_SILU_SPLIT=$sp timeout 300 python benchmarks/bench_layer_glue.py --tokens 256 2>/dev/null | grep silu_and_mul_dynamic >> $S
done
timeout 300 python benchmarks/bench_layer_glue.py --tokens 256 > gpurun_out/r63_glue_256.jsonl 2> gpurun_out/r63_glue_256.err; echo "glue 256 exit $?" >> $S
timeout 300 python benchmarks/bench_layer_glue.py --tokens 64 > gpurun_out/r63_glue_64.jsonl 2> gpurun_out/r63_glue_64.err; echo "glue 64 exit $?" >> $S
cat gpurun_out/r63_glue_256.jsonl gpurun_out/r63_glue_64.jsonl >> $S
cat $S
