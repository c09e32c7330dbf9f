#!/bin/bash
# GPU session 54 (1 GPU): persistent paged-attention kernel: parity with it forced on, then the attention sweep with the heuristic.
mkdir -p gpurun_out; S=gpurun_out/r54_summary.txt; : > $S
VB200_PA_PERSIST=1 timeout 600 python -m pytest tests/test_gpu_attention.py tests/test_gpu_decode_step.py tests/test_gpu_decode_loop.py -m gpu -q --timeout 120 -x > gpurun_out/r54_tests_forced.log 2>&1; echo "forced-persistent tests exit $?" >> $S
tail -5 gpurun_out/r54_tests_forced.log >> $S
timeout 600 python -m pytest tests/test_gpu_attention.py -m gpu -q --timeout 120 -x > gpurun_out/r54_tests_auto.log 2>&1; echo "auto tests exit $?" >> $S
tail -3 gpurun_out/r54_tests_auto.log >> $S
timeout 600 python benchmarks/bench_paged_attention.py --no-ref > gpurun_out/r54_pa_auto.log 2>&1; echo "sweep auto exit $?" >> $S
VB200_PA_PERSIST=0 timeout 600 python benchmarks/bench_paged_attention.py --no-ref > gpurun_out/r54_pa_off.log 2>&1; echo "sweep off exit $?" >> $S
VB200_PA_PERSIST=1 timeout 600 python benchmarks/bench_paged_attention.py --no-ref > gpurun_out/r54_pa_forced.log 2>&1; echo "sweep forced exit $?" >> $S
for f in auto off forced; do echo "== $f" >> $S; grep '^{' gpurun_out/r54_pa_$f.log | python -c "
import sys,json
for l in sys.stdin:
    d=json.loads(l); print(d['cfg'], d['ours_us'], d['ours_GBs'], [v for k,v in d.items() if k.startswith('frac_of')])" >> $S; done
cat $S
