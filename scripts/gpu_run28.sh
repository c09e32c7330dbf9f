#!/bin/bash
# GPU session 28 (1 GPU): ncu --set full of the paged-attention kernel at the config-5 shape (1 kv head / 8 q heads per GPU).
mkdir -p gpurun_out
timeout 120 python benchmarks/bench_paged_attention.py --cfg 5 --no-ref > gpurun_out/r28_pa_cfg5.log 2>&1; echo "plain exit $?"
timeout 300 ncu --set full --clock-control none --import-source on -k regex:paged_attention_tma -s 5 -c 1 -o gpurun_out/r28_pa_cfg5_full python benchmarks/bench_paged_attention.py --cfg 5 --no-ref --iters 4 > gpurun_out/r28_ncu.log 2>&1; echo "ncu exit $?"
grep -h '^{' gpurun_out/r28_pa_cfg5.log | cut -c1-300
