#!/bin/bash
# GPU session 61 (1 GPU): where the e2e step spends its time on the device (events around H2D / graph / D2H).
mkdir -p gpurun_out; S=gpurun_out/r61_summary.txt; : > $S
VB200_E2E_PROBE=1 timeout 600 python bench.py --no-cpu-baseline > gpurun_out/r61_bench_default.log 2>&1; echo "bench default exit $?" >> $S
grep -h '^{' gpurun_out/r61_bench_default.log | python -c "
import sys, json
d = json.loads(sys.stdin.read())
print({k: d[k] for k in ('value', 'ms_per_step', 'e2e', 'gpu_launches_per_step', 'clocks')})" >> $S
grep e2e-probe gpurun_out/r61_bench_default.log >> $S
cat $S
