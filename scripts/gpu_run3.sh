#!/bin/bash
mkdir -p gpurun_out; S=gpurun_out/r3_summary.txt; : > $S
timeout 600 python -m pytest tests/test_gpu_layer_glue.py -m gpu -q --timeout 120 > gpurun_out/r3_glue.log 2>&1; echo "glue exit $?" >> $S
timeout 600 python benchmarks/bench_scaled_mm.py --M 16,128,256,1024 > gpurun_out/r3_bench_mm.log 2>&1; echo "bench_mm exit $?" >> $S
CMD="python bench.py --steps 3 --warmup 3 --no-cpu-baseline --profile-window"
timeout 600 $CMD > gpurun_out/r3_plain.log 2>&1 &&
timeout 900 ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none -c 420 --csv --log-file gpurun_out/r3_launches.csv $CMD > gpurun_out/r3_ncu_launch.log 2>&1
echo "ncu launches exit $?" >> $S
cat $S
