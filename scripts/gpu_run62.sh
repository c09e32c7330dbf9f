#!/bin/bash
# GPU session 62 (1 GPU): variable-length prefill over the paged flash-layout cache: parity + speed vs the library kernel.
mkdir -p gpurun_out; S=gpurun_out/r62_summary.txt; : > $S
timeout 900 python -m pytest tests/test_gpu_flash_attention.py tests/test_c_abi.py -q --timeout 300 > gpurun_out/r62_tests.log 2>&1; echo "tests exit $?" >> $S
tail -15 gpurun_out/r62_tests.log | cut -c1-300 >> $S
timeout 300 python benchmarks/bench_prefill.py > gpurun_out/r62_prefill.jsonl 2> gpurun_out/r62_prefill.err; echo "bench prefill exit $?" >> $S
cat gpurun_out/r62_prefill.jsonl >> $S; tail -5 gpurun_out/r62_prefill.err >> $S
cat $S
