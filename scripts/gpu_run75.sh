#!/bin/bash
# GPU session 75 (1 GPU): fp8 MoE pipeline with per-tensor scales: quantise while gathering, abs-max fused with the
# activation (8 launches per layer instead of 12): bit-exact tests, MoE suite, config-4 bench + launch list.
mkdir -p gpurun_out; S=gpurun_out/r75_summary.txt; : > $S
timeout 400 python -m pytest tests/test_gpu_fused_moe.py tests/test_gpu_cutlass_moe.py tests/test_moe_layer.py tests/test_gpu_quant.py tests/test_c_abi.py tests/test_gpu_decode_step.py -m gpu -q --timeout 120 > gpurun_out/r75_tests_moe.log 2>&1; echo "moe tests exit $?" >> $S
tail -4 gpurun_out/r75_tests_moe.log >> $S
grep -E "^FAILED|^ERROR" gpurun_out/r75_tests_moe.log | head -20 >> $S
timeout 600 python bench.py --config cfg4 --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/r75_bench_cfg4.log 2>&1; echo "bench cfg4 exit $?" >> $S
grep -h '^{' gpurun_out/r75_bench_cfg4.log | python -c "
import sys, json
d = json.loads(sys.stdin.read())
print({k: d[k] for k in ('value', 'ms_per_step', 'gpu_launches_per_step', 'clocks')})" >> $S
CMD="python bench.py --config cfg4 --steps 2 --warmup 3 --no-cpu-baseline --profile-window"
timeout 600 ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none -c 1400 --csv --log-file gpurun_out/r75_launches_cfg4.csv $CMD > gpurun_out/r75_ncu_launch_cfg4.log 2>&1; echo "ncu cfg4 launches exit $?" >> $S
python scripts/agg_launches.py gpurun_out/r75_launches_cfg4.csv | head -28 | cut -c1-140 >> $S
cat $S
