#!/bin/bash
# GPU session 9: split-K reduce through vectorised partials + bulk fetch; GEMM ncu captures; bench + launch list.
mkdir -p gpurun_out; S=gpurun_out/r9_summary.txt; : > $S
timeout 600 python -m pytest tests/test_gpu_scaled_mm.py -m gpu -q --timeout 60 -x > gpurun_out/r9_mm.log 2>&1; echo "mm exit $?" >> $S
tail -5 gpurun_out/r9_mm.log >> $S
timeout 300 python benchmarks/bench_scaled_mm.py --M 16,64,128,256,1024,4096 > gpurun_out/r9_bench_mm.log 2>&1; echo "bench_mm exit $?" >> $S
timeout 900 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/r9_bench_cfg3.log 2>&1; echo "bench exit $?" >> $S
CMD="python benchmarks/prof_scaled_mm.py o,gate_up 256"
timeout 300 $CMD > gpurun_out/r9_prof_plain.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:scaled_mm_kernel -s 4 -c 1 -o gpurun_out/r9_mm_o256 $CMD > gpurun_out/r9_ncu1.log 2>&1
echo "ncu o exit $?" >> $S
timeout 900 ncu --set full --clock-control none --import-source on -k regex:scaled_mm_kernel -s 10 -c 1 -o gpurun_out/r9_mm_gu256 $CMD > gpurun_out/r9_ncu2.log 2>&1
echo "ncu gu exit $?" >> $S
CMD="python benchmarks/prof_scaled_mm.py gate_up 4096"
timeout 300 $CMD > gpurun_out/r9_prof_plain2.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:scaled_mm_kernel -s 4 -c 1 -o gpurun_out/r9_mm_gu4096 $CMD > gpurun_out/r9_ncu3.log 2>&1
echo "ncu gu4096 exit $?" >> $S
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --profile-window"
timeout 600 $CMD > gpurun_out/r9_plain.log 2>&1 &&
timeout 900 ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none -c 800 --csv --log-file gpurun_out/r9_launches.csv $CMD > gpurun_out/r9_ncu_launch.log 2>&1
echo "ncu launches exit $?" >> $S
cat $S
