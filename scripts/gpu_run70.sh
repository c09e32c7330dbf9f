#!/bin/bash
# GPU session 70 (1 GPU): log-sum-exp output + cascade attention + head size 192 in the variable-length kernel.
mkdir -p gpurun_out; S=gpurun_out/r70_summary.txt; : > $S
timeout 900 python -m pytest tests/test_gpu_flash_attention.py tests/test_c_abi.py -m gpu -q --timeout 300 -k "cascade or lse or rejects or golden or c_abi or backend" > gpurun_out/r70_tests_cascade.log 2>&1; echo "cascade tests exit $?" >> $S
tail -4 gpurun_out/r70_tests_cascade.log >> $S
grep -E "^FAILED|^ERROR" gpurun_out/r70_tests_cascade.log | head -30 >> $S
timeout 900 python -m pytest tests/test_gpu_flash_attention.py -m gpu -q --timeout 300 -x > gpurun_out/r70_tests_flash.log 2>&1; echo "flash tests exit $?" >> $S
tail -3 gpurun_out/r70_tests_flash.log >> $S
cat $S
