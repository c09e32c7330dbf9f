#!/bin/bash
# GPU session 10: HBM read ceiling probe; paged-attention partition sweep and L2-prefetch distance sweep.
mkdir -p gpurun_out; S=gpurun_out/r10_summary.txt; : > $S
timeout 120 benchmarks/bin/hbm_read_probe > gpurun_out/r10_read_probe.log 2>&1; echo "probe exit $?" >> $S
timeout 300 python -m pytest tests/test_gpu_attention.py -m gpu -q --timeout 120 -x > gpurun_out/r10_pa_tests.log 2>&1; echo "pa tests (pf 0) exit $?" >> $S
VB200_PA_PREFETCH=4 timeout 300 python -m pytest tests/test_gpu_attention.py -m gpu -q --timeout 120 -x > gpurun_out/r10_pa_tests_pf4.log 2>&1; echo "pa tests (pf 4) exit $?" >> $S
for pf in 0 2 4 8 16; do
  VB200_PA_PREFETCH=$pf timeout 300 python benchmarks/bench_paged_attention.py --no-ref --iters 30 > gpurun_out/r10_pa_pf$pf.log 2>&1; echo "pf $pf exit $?" >> $S
done
for part in 4096 2048 1376 1024 832 704 512; do
  timeout 300 python benchmarks/bench_paged_attention.py --no-ref --iters 30 --cfg 3 --partition $part > gpurun_out/r10_pa_part$part.log 2>&1; echo "part $part exit $?" >> $S
done
for part in 8192 4096 2048 1024 512; do
  timeout 300 python benchmarks/bench_paged_attention.py --no-ref --iters 30 --cfg 5 --partition $part > gpurun_out/r10_pa5_part$part.log 2>&1
done
for part in 2048 1024 704 512 256; do
  timeout 300 python benchmarks/bench_paged_attention.py --no-ref --iters 30 --cfg 2 --partition $part > gpurun_out/r10_pa2_part$part.log 2>&1
done
cat $S
