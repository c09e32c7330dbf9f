#!/bin/bash
# GPU session 80a (1 GPU): soft-cap tanh, exact two-MUFU form (shipped) vs tanh.approx (alternate build): parity + timing.
mkdir -p gpurun_out; S=gpurun_out/r80a_summary.txt; : > $S
ALT=$PWD/benchmarks/bin/libvllm_b200_tanhapprox.so
VLLM_B200_LIB=$ALT timeout 400 python -m pytest tests/test_gpu_flash_attention.py -m gpu -q --timeout 120 -k "softcap or cascade or alibi" > gpurun_out/r80a_tests_approx.log 2>&1; echo "soft-cap tests with tanh.approx exit $?" >> $S
tail -3 gpurun_out/r80a_tests_approx.log >> $S
grep -E "^FAILED" gpurun_out/r80a_tests_approx.log | wc -l >> $S
echo "## shipped (1 - 2 / (2^(2x log2 e) + 1))" >> $S
timeout 300 python benchmarks/bench_prefill.py 2> /dev/null | grep -i "cap" >> $S
echo "## tanh.approx" >> $S
VLLM_B200_LIB=$ALT timeout 300 python benchmarks/bench_prefill.py 2> /dev/null | grep -i "cap" >> $S
cat $S
