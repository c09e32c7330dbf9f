#!/bin/bash
# GPU session 8: CTA-pair (cta_group::2) persistent GEMM: parity first (under a short timeout: a wrong barrier
# protocol hangs), then microbench new vs old kernel, then the whole suite and the bench.
mkdir -p gpurun_out; S=gpurun_out/r8_summary.txt; : > $S
timeout 120 python -m pytest tests/test_gpu_scaled_mm.py -m gpu -q --timeout 60 -x -k "shapes and 128-128" > gpurun_out/r8_mm_first.log 2>&1; echo "mm first exit $?" >> $S
tail -3 gpurun_out/r8_mm_first.log >> $S
timeout 600 python -m pytest tests/test_gpu_scaled_mm.py -m gpu -q --timeout 60 > gpurun_out/r8_mm.log 2>&1; echo "mm exit $?" >> $S
tail -15 gpurun_out/r8_mm.log >> $S
timeout 300 python benchmarks/bench_scaled_mm.py --M 16,64,128,256,1024,4096 > gpurun_out/r8_bench_mm_v2.log 2>&1; echo "bench_mm v2 exit $?" >> $S
timeout 900 python -m pytest tests -m gpu -q --timeout 300 > gpurun_out/r8_tests.log 2>&1; echo "tests exit $?" >> $S
tail -5 gpurun_out/r8_tests.log >> $S
timeout 900 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/r8_bench_cfg3.log 2>&1; echo "bench exit $?" >> $S
cat $S
