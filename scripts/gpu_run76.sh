#!/bin/bash
# GPU session 76 (1 GPU): unquantised dense linear (lm_head, bf16 projections) on the tcgen05 grouped GEMM vs the library GEMM.
mkdir -p gpurun_out; S=gpurun_out/r76_summary.txt; : > $S
timeout 400 python -m pytest tests/test_gpu_dense_linear.py tests/test_gpu_decode_step.py tests/test_gpu_decode_loop.py -m gpu -q --timeout 120 > gpurun_out/r76_tests.log 2>&1; echo "tests exit $?" >> $S
tail -4 gpurun_out/r76_tests.log >> $S
grep -E "^FAILED|^ERROR" gpurun_out/r76_tests.log | head -20 >> $S
for D in 0 1; do
  for CFG in cfg3 cfg2; do
    VB200_DENSE16=$D timeout 600 python bench.py --config $CFG --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/r76_bench_${CFG}_dense$D.log 2>&1; echo "bench $CFG dense16=$D exit $?" >> $S
    grep -h '^{' gpurun_out/r76_bench_${CFG}_dense$D.log | python -c "
import sys, json
d = json.loads(sys.stdin.read())
print({k: d[k] for k in ('value', 'ms_per_step', 'gpu_launches_per_step')}, d['e2e']['value'], d['clocks']['sm_mhz'])" >> $S
  done
done
cat $S
