#!/bin/bash
# GPU session 86 (1 GPU): the remaining suites that reach the attention entry points after the decode-window change.
mkdir -p gpurun_out; S=gpurun_out/r86_summary.txt; : > $S
timeout 150 python -m pytest tests/test_gpu_torch_ops.py tests/test_gpu_dropin.py tests/test_gpu_decode_loop.py tests/test_gpu_cache.py -m gpu -q --timeout 100 > gpurun_out/r86_tests_a.log 2>&1; echo "torch ops / drop-in / decode loop / cache exit $?" >> $S
tail -2 gpurun_out/r86_tests_a.log >> $S
timeout 100 python -m pytest tests/test_gpu_flash_attention.py -m gpu -q --timeout 100 -k "cascade or alibi or lse or golden or fp8_cache" > gpurun_out/r86_tests_b.log 2>&1; echo "flash cascade / alibi / lse / golden / fp8 exit $?" >> $S
tail -2 gpurun_out/r86_tests_b.log >> $S
cat $S
