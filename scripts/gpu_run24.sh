#!/bin/bash
# GPU session 24 (1 GPU): final state: full parity suite, smoke(), default bench + reference arm, attention microbench.
mkdir -p gpurun_out; S=gpurun_out/r24_summary.txt; : > $S
timeout 1200 python -m pytest tests -m gpu -q --timeout 180 > gpurun_out/r24_tests.log 2>&1; echo "tests exit $?" >> $S
tail -4 gpurun_out/r24_tests.log >> $S
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r24_smoke.log 2>&1; echo "smoke exit $?" >> $S
tail -1 gpurun_out/r24_smoke.log >> $S
timeout 200 python benchmarks/bench_paged_attention.py > gpurun_out/r24_bench_pa.log 2>&1; echo "bench_pa exit $?" >> $S
timeout 600 python bench.py > gpurun_out/r24_bench_default.log 2>&1; echo "bench default exit $?" >> $S
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r24_bench_ref.log 2>&1; echo "bench ref exit $?" >> $S
grep -h '^{' gpurun_out/r24_bench_default.log gpurun_out/r24_bench_ref.log | cut -c1-300 >> $S
cat $S
