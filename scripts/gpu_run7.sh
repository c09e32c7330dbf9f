#!/bin/bash
# GPU session 7 (re-entry): whole gpu suite, GEMM microbench after the cluster split-K change, bench both arms, launch list.
mkdir -p gpurun_out; S=gpurun_out/r7_summary.txt; : > $S
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv >> $S
timeout 1500 python -m pytest tests -m gpu -q --timeout 300 > gpurun_out/r7_tests.log 2>&1; echo "tests exit $?" >> $S
tail -5 gpurun_out/r7_tests.log >> $S
timeout 600 python benchmarks/bench_scaled_mm.py --M 16,64,128,256,1024 > gpurun_out/r7_bench_mm.log 2>&1; echo "bench_mm exit $?" >> $S
timeout 600 python benchmarks/bench_paged_attention.py > gpurun_out/r7_bench_pa.log 2>&1; echo "bench_pa exit $?" >> $S
timeout 900 python bench.py --steps 10 --warmup 3 > gpurun_out/r7_bench_cfg3.log 2>&1; echo "bench exit $?" >> $S
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r7_bench_ref.log 2>&1; echo "bench ref exit $?" >> $S
CMD="python bench.py --steps 3 --warmup 3 --no-cpu-baseline"
timeout 600 $CMD > gpurun_out/r7_plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -s 450 -c 460 --csv --log-file gpurun_out/r7_launches.csv $CMD > gpurun_out/r7_ncu_launch.log 2>&1
echo "ncu launches exit $?" >> $S
cat $S
