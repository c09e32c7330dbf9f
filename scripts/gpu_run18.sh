#!/bin/bash
# GPU session 18 (1 GPU): decode-loop kernels parity, GEMM phase trace, cfg5 attention partition sweep.
mkdir -p gpurun_out; S=gpurun_out/r18_summary.txt; : > $S
timeout 300 python -m pytest tests/test_gpu_decode_loop.py tests/test_gpu_torch_ops.py -m gpu -q --timeout 60 > gpurun_out/r18_new_tests.log 2>&1; echo "new tests exit $?" >> $S
tail -12 gpurun_out/r18_new_tests.log >> $S
timeout 120 python benchmarks/trace_scaled_mm.py > gpurun_out/r18_mm_trace.log 2>&1; echo "trace exit $?" >> $S
for P in 0 128 160 224 256 320 448 512 704 1024 2048 4096; do
  echo "partition $P" >> gpurun_out/r18_pa_sweep.log
  timeout 60 python benchmarks/bench_paged_attention.py --cfg 5 --no-ref --partition $P >> gpurun_out/r18_pa_sweep.log 2>&1
done
echo "sweep done" >> $S
cat $S
