#!/bin/bash
# GPU session 26 (1 GPU): Mixtral (cfg4) launch list of the timed window + ncu --set full of the grouped MoE GEMM.
mkdir -p gpurun_out; S=gpurun_out/r26_summary.txt; : > $S
CMD="python bench.py --config cfg4 --steps 2 --warmup 3 --no-cpu-baseline --profile-window"
timeout 400 $CMD > gpurun_out/r26_bench_cfg4_plain.log 2>&1; echo "plain run exit $?" >> $S
timeout 500 ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none -c 1400 --csv --log-file gpurun_out/r26_launches_cfg4.csv $CMD > gpurun_out/r26_ncu_launch.log 2>&1; echo "ncu launches exit $?" >> $S
timeout 500 ncu --profile-from-start off --set full --clock-control none --import-source on -k regex:moe_grouped_gemm -c 2 -o gpurun_out/r26_moe_full $CMD > gpurun_out/r26_ncu_moe.log 2>&1; echo "ncu moe exit $?" >> $S
cat $S
