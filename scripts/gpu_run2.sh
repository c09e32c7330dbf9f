#!/bin/bash
# GPU session 2: new glue tests, first bench.py lines, ncu launch list + full capture of the dominant kernel.
mkdir -p gpurun_out; S=gpurun_out/r2_summary.txt; : > $S
timeout 600 python -m pytest tests/test_gpu_layer_glue.py -m gpu -q --timeout 120 > gpurun_out/r2_glue.log 2>&1; echo "glue exit $?" >> $S
timeout 300 python bench.py --config tiny --steps 5 --warmup 3 > gpurun_out/r2_bench_tiny.log 2>&1; echo "bench tiny exit $?" >> $S
timeout 900 python bench.py --steps 10 --warmup 3 > gpurun_out/r2_bench_cfg3.log 2>&1; echo "bench cfg3 exit $?" >> $S
CMD="python bench.py --steps 3 --warmup 3 --no-cpu-baseline"
timeout 600 $CMD > gpurun_out/r2_plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -s 450 -c 460 --csv --log-file gpurun_out/r2_launches.csv $CMD > gpurun_out/r2_ncu_launch.log 2>&1
echo "ncu launches exit $?" >> $S
timeout 900 ncu --set full --clock-control none --import-source on -k regex:paged_attention_kernel -s 40 -c 2 -o gpurun_out/r2_pa_full $CMD > gpurun_out/r2_ncu_full.log 2>&1
echo "ncu full exit $?" >> $S
cat $S
