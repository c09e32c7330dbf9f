#!/bin/bash
# GPU session 14: GEMM epilogue (pipelined tcgen05.ld, explicit ld.shared, self re-arming split-K counter) +
# bulk-copy attention: parity (short timeouts: a wrong barrier protocol hangs), microbenchmarks, trace, bench.
mkdir -p gpurun_out; S=gpurun_out/r14_summary.txt; : > $S
timeout 240 python -m pytest tests/test_gpu_scaled_mm.py -m gpu -q --timeout 60 -x > gpurun_out/r14_mm_tests.log 2>&1; rc=$?; echo "mm tests exit $rc" >> $S
tail -3 gpurun_out/r14_mm_tests.log >> $S
if [ $rc -ne 0 ]; then cat $S; exit 1; fi
timeout 600 python -m pytest tests -m gpu -q --timeout 120 -x > gpurun_out/r14_tests.log 2>&1; echo "tests exit $?" >> $S
tail -4 gpurun_out/r14_tests.log >> $S
timeout 200 python benchmarks/bench_scaled_mm.py --M 16,64,128,256,1024,4096 > gpurun_out/r14_bench_mm.log 2>&1; echo "bench_mm exit $?" >> $S
timeout 100 python benchmarks/trace_scaled_mm.py > gpurun_out/r14_mm_trace.log 2>&1; echo "trace exit $?" >> $S
timeout 300 python bench.py --steps 10 --warmup 3 > gpurun_out/r14_bench_cfg3.log 2>&1; echo "bench exit $?" >> $S
timeout 200 python bench.py --config cfg2 --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/r14_bench_cfg2.log 2>&1; echo "bench cfg2 exit $?" >> $S
cat $S
