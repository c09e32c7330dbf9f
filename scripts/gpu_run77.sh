#!/bin/bash
# GPU session 77 (1 GPU): MoE router in one launch (gate GEMM + fused_topk + abs-max of the hidden states).
mkdir -p gpurun_out; S=gpurun_out/r77_summary.txt; : > $S
timeout 400 python -m pytest tests/test_gpu_moe_routing.py tests/test_gpu_fused_moe.py tests/test_gpu_decode_step.py tests/test_c_abi.py -m gpu -q --timeout 120 > gpurun_out/r77_tests.log 2>&1; echo "tests exit $?" >> $S
tail -4 gpurun_out/r77_tests.log >> $S
grep -E "^FAILED|^ERROR" gpurun_out/r77_tests.log | head -20 >> $S
for R in 0 1; do
  VB200_FUSED_ROUTER=$R timeout 600 python bench.py --config cfg4 --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/r77_bench_cfg4_router$R.log 2>&1; echo "bench cfg4 fused router=$R exit $?" >> $S
  grep -h '^{' gpurun_out/r77_bench_cfg4_router$R.log | python -c "
import sys, json
d = json.loads(sys.stdin.read())
print({k: d[k] for k in ('value', 'ms_per_step', 'gpu_launches_per_step')}, d['e2e']['value'], d['clocks']['sm_mhz'])" >> $S
done
cat $S
