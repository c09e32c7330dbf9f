#!/bin/bash
# GPU session 87 (1 GPU): logits soft-cap on the flash-layout decode kernel: parity.
mkdir -p gpurun_out; S=gpurun_out/r87_summary.txt; : > $S
timeout 120 python -m pytest tests/test_gpu_flash_attention.py -m gpu -q --timeout 100 -k "decode_softcap or decode_with_sliding_window or backend or rejects or (window_softcap and n3q1 and not n3q129) or paged_attention_flash" > gpurun_out/r87_tests.log 2>&1; echo "decode soft-cap / window tests exit $?" >> $S
tail -3 gpurun_out/r87_tests.log >> $S
grep -E "^FAILED|^ERROR" gpurun_out/r87_tests.log | head >> $S
cat $S
