#!/bin/bash
mkdir -p gpurun_out; S=gpurun_out/r5_summary.txt; : > $S
timeout 600 python -m pytest tests/test_gpu_scaled_mm.py -m gpu -q --timeout 120 -x > gpurun_out/r5_mm.log 2>&1; echo "mm exit $?" >> $S
timeout 600 python -m pytest tests/test_gpu_torch_ops.py tests/test_gpu_decode_step.py -m gpu -q --timeout 300 > gpurun_out/r5_ops.log 2>&1; echo "ops exit $?" >> $S
timeout 600 python benchmarks/bench_scaled_mm.py --M 16,64,128,256 > gpurun_out/r5_bench_mm.log 2>&1; echo "bench_mm exit $?" >> $S
timeout 900 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/r5_bench_cfg3.log 2>&1; echo "bench exit $?" >> $S
cat $S
