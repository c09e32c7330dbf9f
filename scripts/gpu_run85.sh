#!/bin/bash
# GPU session 85 (1 GPU): sliding window on the decode kernel: parity (new tests, the decode-only window cases, the
# paged-layout suite whose reduce kernel gained an argument, the decode step) and timing vs the variable-length kernel.
mkdir -p gpurun_out; S=gpurun_out/r85_summary.txt; : > $S
timeout 300 python -m pytest tests/test_gpu_flash_attention.py -m gpu -q --timeout 120 -k "decode_with_sliding_window or backend or rejects or (window_softcap and n3q1 and not n3q129) or paged_attention_flash" > gpurun_out/r85_tests_flash.log 2>&1; echo "flash window tests exit $?" >> $S
tail -3 gpurun_out/r85_tests_flash.log >> $S
grep -E "^FAILED|^ERROR" gpurun_out/r85_tests_flash.log | head >> $S
timeout 300 python -m pytest tests/test_gpu_attention.py tests/test_gpu_decode_step.py tests/test_c_abi.py tests/test_zz_no_alloc.py -m gpu -q --timeout 120 > gpurun_out/r85_tests_paged.log 2>&1; echo "paged-layout tests exit $?" >> $S
tail -2 gpurun_out/r85_tests_paged.log >> $S
timeout 200 python benchmarks/bench_window_decode.py > gpurun_out/r85_window_decode.jsonl 2> gpurun_out/r85_window_decode.err; echo "bench exit $?" >> $S
cat gpurun_out/r85_window_decode.jsonl >> $S
cat $S
