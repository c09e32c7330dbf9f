#!/bin/bash
# GPU session 69 (1 GPU): sliding window + logits soft-cap in the variable-length attention kernel: parity tests, the
# tests that need the reference CUTLASS build, prefill benchmark incl. window / soft-cap cases vs the library kernel.
mkdir -p gpurun_out; S=gpurun_out/r69_summary.txt; : > $S
timeout 900 python -m pytest tests/test_gpu_flash_attention.py tests/test_c_abi.py -m gpu -q --timeout 300 > gpurun_out/r69_tests_flash.log 2>&1; echo "flash tests exit $?" >> $S
tail -4 gpurun_out/r69_tests_flash.log >> $S
grep -E "^FAILED|^ERROR" gpurun_out/r69_tests_flash.log | head -30 >> $S
timeout 600 python -m pytest tests/test_gpu_scaled_mm.py tests/test_gpu_cutlass_moe.py tests/test_gpu_fused_moe.py tests/test_gpu_dropin.py -m gpu -q --timeout 300 > gpurun_out/r69_tests_cutlass.log 2>&1; echo "cutlass-ref tests exit $?" >> $S
tail -3 gpurun_out/r69_tests_cutlass.log >> $S
timeout 300 python benchmarks/bench_prefill.py > gpurun_out/r69_prefill.jsonl 2> gpurun_out/r69_prefill.err; echo "bench_prefill exit $?" >> $S
cat gpurun_out/r69_prefill.jsonl >> $S
cat $S
