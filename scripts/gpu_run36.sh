#!/bin/bash
# GPU session 36 (round 2, 1 GPU): the restored round-2 tree + rewritten cutlass_moe: full parity suite, smoke, default bench,
# GEMM sweep vs cuBLASLt, launch list of the timed window, GEMM per-phase trace + plan sweep.
mkdir -p gpurun_out; S=gpurun_out/r36_summary.txt; : > $S
timeout 1500 python -m pytest tests -m gpu -q --timeout 300 > gpurun_out/r36_tests.log 2>&1; echo "tests exit $?" >> $S
tail -6 gpurun_out/r36_tests.log >> $S
grep -E "^FAILED|^ERROR" gpurun_out/r36_tests.log | head -20 >> $S
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r36_smoke.log 2>&1; echo "smoke exit $?" >> $S
timeout 600 python bench.py --no-cpu-baseline > gpurun_out/r36_bench_default.log 2>&1; echo "bench default exit $?" >> $S
timeout 600 python benchmarks/bench_scaled_mm.py > gpurun_out/r36_bench_mm.log 2>&1; echo "bench_mm exit $?" >> $S
timeout 300 python benchmarks/trace_scaled_mm.py > gpurun_out/r36_mm_trace.log 2>&1; echo "trace exit $?" >> $S
timeout 900 python benchmarks/sweep_scaled_mm_plan.py --M 16,128,256 > gpurun_out/r36_plan_sweep.log 2>&1; echo "plan sweep exit $?" >> $S
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --profile-window"
timeout 600 ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none -c 900 --csv --log-file gpurun_out/r36_launches.csv $CMD > gpurun_out/r36_ncu_launch.log 2>&1; echo "ncu launches exit $?" >> $S
grep -h '^{' gpurun_out/r36_bench_default.log | cut -c1-600 >> $S
cat $S
