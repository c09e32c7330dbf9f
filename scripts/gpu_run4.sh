#!/bin/bash
mkdir -p gpurun_out; S=gpurun_out/r4_summary.txt; : > $S
timeout 900 python -m pytest tests/test_gpu_fused_moe.py -m gpu -q --timeout 180 -x > gpurun_out/r4_moe.log 2>&1; echo "moe exit $?" >> $S
timeout 900 python -m pytest tests/test_gpu_decode_step.py -m gpu -q --timeout 300 > gpurun_out/r4_decode.log 2>&1; echo "decode exit $?" >> $S
cat $S
