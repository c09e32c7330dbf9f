#!/bin/bash
# GPU session 68 (1 GPU): no-allocation attention entry point (alone in a fresh process, then after the whole suite),
# full parity suite, smoke, default bench + reference arm.
mkdir -p gpurun_out; S=gpurun_out/r68_summary.txt; : > $S
timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -q --timeout 200 > gpurun_out/r68_tests_no_alloc_alone.log 2>&1; echo "no_alloc alone exit $?" >> $S
tail -3 gpurun_out/r68_tests_no_alloc_alone.log >> $S
timeout 1500 python -m pytest tests -m gpu -x -q --timeout 300 > gpurun_out/r68_tests.log 2>&1; echo "tests exit $?" >> $S
tail -6 gpurun_out/r68_tests.log >> $S
grep -E "^FAILED|^ERROR" gpurun_out/r68_tests.log | head -20 >> $S
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r68_smoke.log 2>&1; echo "smoke exit $?" >> $S
timeout 600 python bench.py > gpurun_out/r68_bench_default.log 2>&1; echo "bench default exit $?" >> $S
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r68_bench_reference.log 2>&1; echo "bench reference exit $?" >> $S
grep -h '^{' gpurun_out/r68_bench_default.log gpurun_out/r68_bench_reference.log | cut -c1-400 >> $S
cat $S
