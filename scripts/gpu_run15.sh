#!/bin/bash
# GPU session 15: re-establish state after container re-creation: full parity suite, GEMM microbench,
# bench cfg3 (+ reference arm), ncu launch list of the bench command.
mkdir -p gpurun_out; S=gpurun_out/r15_summary.txt; : > $S
timeout 900 python -m pytest tests -m gpu -q --timeout 120 -x > gpurun_out/r15_tests.log 2>&1; echo "tests exit $?" >> $S
tail -4 gpurun_out/r15_tests.log >> $S
timeout 200 python benchmarks/bench_scaled_mm.py --M 16,64,128,256,1024,4096 > gpurun_out/r15_bench_mm.log 2>&1; echo "bench_mm exit $?" >> $S
timeout 200 python benchmarks/bench_paged_attention.py > gpurun_out/r15_bench_pa.log 2>&1; echo "bench_pa exit $?" >> $S
timeout 400 python bench.py --steps 10 --warmup 3 > gpurun_out/r15_bench_cfg3.log 2>&1; echo "bench exit $?" >> $S
timeout 300 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r15_bench_ref.log 2>&1; echo "bench ref exit $?" >> $S
timeout 200 python bench.py --config cfg2 --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/r15_bench_cfg2.log 2>&1; echo "bench cfg2 exit $?" >> $S
timeout 500 ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file gpurun_out/r15_launches.csv python bench.py --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/r15_ncu.log 2>&1; echo "ncu launches exit $?" >> $S
cat $S
