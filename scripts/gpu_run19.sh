#!/bin/bash
# GPU session 19 (1 GPU): GEMM token-split plan (parity + sweep), cfg5 attention with consecutive physical blocks
# (DRAM-locality hypothesis), bench cfg3.
mkdir -p gpurun_out; S=gpurun_out/r19_summary.txt; : > $S
timeout 400 python -m pytest tests/test_gpu_scaled_mm.py -m gpu -q --timeout 60 -x > gpurun_out/r19_mm_tests.log 2>&1; echo "mm tests exit $?" >> $S
tail -3 gpurun_out/r19_mm_tests.log >> $S
timeout 200 python benchmarks/bench_scaled_mm.py --M 128,192,256 > gpurun_out/r19_bench_mm.log 2>&1; echo "bench_mm exit $?" >> $S
timeout 100 python benchmarks/trace_scaled_mm.py > gpurun_out/r19_mm_trace.log 2>&1; echo "trace exit $?" >> $S
PA_SEQ_BLOCKS=1 timeout 100 python benchmarks/bench_paged_attention.py --cfg 5 --no-ref > gpurun_out/r19_pa_cfg5_seqblocks.log 2>&1
PA_SEQ_BLOCKS=1 timeout 100 python benchmarks/bench_paged_attention.py --cfg 5 --no-ref --partition 512 >> gpurun_out/r19_pa_cfg5_seqblocks.log 2>&1
timeout 400 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/r19_bench_cfg3.log 2>&1; echo "bench exit $?" >> $S
cat $S; grep -h '^{' gpurun_out/r19_pa_cfg5_seqblocks.log | cut -c1-300
