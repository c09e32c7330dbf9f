#!/bin/bash
# GPU session 35 (4 epilogue warps again, relaxed teardown barrier, stream-K): the cluster / DSMEM split-K GEMM with weight prefetch ahead of griddepcontrol.wait.
mkdir -p gpurun_out; S=gpurun_out/r35_summary.txt; : > $S
timeout 900 python -m pytest tests/test_gpu_scaled_mm.py tests/test_gpu_decode_step.py tests/test_gpu_dropin.py -m gpu -q --timeout 300 -x > gpurun_out/r35_tests.log 2>&1; echo "mm+decode tests exit $?" >> $S
tail -8 gpurun_out/r35_tests.log >> $S
timeout 900 python benchmarks/sweep_scaled_mm_plan.py > gpurun_out/r35_plan_sweep.log 2>&1; echo "plan sweep exit $?" >> $S
timeout 600 python benchmarks/bench_scaled_mm.py > gpurun_out/r35_bench_mm.log 2>&1; echo "bench_mm exit $?" >> $S
timeout 300 python benchmarks/trace_scaled_mm.py > gpurun_out/r35_mm_trace.log 2>&1; echo "trace exit $?" >> $S
timeout 600 python bench.py --no-cpu-baseline > gpurun_out/r35_bench_default.log 2>&1; echo "bench default exit $?" >> $S
grep -h '^{' gpurun_out/r35_bench_default.log | cut -c1-300 >> $S
cat $S
