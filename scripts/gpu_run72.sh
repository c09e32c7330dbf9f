#!/bin/bash
# GPU session 72 (1 GPU): grouped MoE GEMM launch order (all M-blocks of a weight-row tile back to back) A/B.
mkdir -p gpurun_out; S=gpurun_out/r72_summary.txt; : > $S
for ORDER in 0 1; do
  VB200_MOE_ORDER=$ORDER timeout 400 python benchmarks/bench_moe.py --tps 1,8 --batches 16,128,512,2048 > gpurun_out/r72_moe_order$ORDER.jsonl 2> gpurun_out/r72_moe_order$ORDER.err; echo "bench_moe order $ORDER exit $?" >> $S
  cut -c1-330 gpurun_out/r72_moe_order$ORDER.jsonl >> $S
done
timeout 600 python -m pytest tests/test_gpu_fused_moe.py tests/test_gpu_cutlass_moe.py tests/test_moe_layer.py -m gpu -q --timeout 300 > gpurun_out/r72_tests_moe.log 2>&1; echo "moe tests exit $?" >> $S
tail -3 gpurun_out/r72_tests_moe.log >> $S
cat $S
