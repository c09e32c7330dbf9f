#!/bin/bash
# GPU session 66 (1 GPU): attention reduce + quant with heads walked together (one wave); launch list + bench.
mkdir -p gpurun_out; S=gpurun_out/r66_summary.txt; : > $S
timeout 900 python -m pytest tests/test_gpu_attention.py tests/test_gpu_decode_step.py -m gpu -q --timeout 300 -x > gpurun_out/r66_tests.log 2>&1; echo "tests exit $?" >> $S
tail -4 gpurun_out/r66_tests.log >> $S
timeout 600 python bench.py --no-cpu-baseline > gpurun_out/r66_bench_default.log 2>&1; echo "bench default exit $?" >> $S
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --profile-window"
timeout 600 ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none -c 900 --csv --log-file gpurun_out/r66_launches.csv $CMD > gpurun_out/r66_ncu_launch.log 2>&1; echo "ncu cfg3 launches exit $?" >> $S
python scripts/agg_launches.py gpurun_out/r66_launches.csv | head -9 >> $S
grep -h '^{' gpurun_out/r66_bench_default.log | python -c "
import sys, json
d = json.loads(sys.stdin.read())
print({k: d[k] for k in ('value', 'ms_per_step', 'gpu_launches_per_step', 'clocks')})
print(d['roofline']['avg_launch_us'], d['roofline']['frac'], d['roofline_gemm']['sum_us_per_layer'])" >> $S
cat $S
