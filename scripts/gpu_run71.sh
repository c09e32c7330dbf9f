#!/bin/bash
# GPU session 71 (1 GPU): ALiBi in the variable-length kernel, BASELINE-sized attention parity, bench_moe with signed
# inputs / relative error.
mkdir -p gpurun_out; S=gpurun_out/r71_summary.txt; : > $S
timeout 600 python -m pytest tests/test_gpu_flash_attention.py tests/test_c_abi.py -m gpu -q --timeout 300 -k "alibi or rejects or c_abi or cascade_backend or golden" > gpurun_out/r71_tests_alibi.log 2>&1; echo "alibi tests exit $?" >> $S
tail -4 gpurun_out/r71_tests_alibi.log >> $S
grep -E "^FAILED|^ERROR" gpurun_out/r71_tests_alibi.log | head -30 >> $S
timeout 900 python -m pytest tests/test_gpu_attention.py -m gpu -q --timeout 600 --durations=5 > gpurun_out/r71_tests_attention.log 2>&1; echo "attention tests exit $?" >> $S
tail -12 gpurun_out/r71_tests_attention.log >> $S
timeout 300 python benchmarks/bench_moe.py --tps 1 --batches 128,512 > gpurun_out/r71_moe.jsonl 2> gpurun_out/r71_moe.err; echo "bench_moe exit $?" >> $S
cat gpurun_out/r71_moe.jsonl >> $S
cat $S
