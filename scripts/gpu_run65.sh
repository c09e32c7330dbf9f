#!/bin/bash
# GPU session 65 (1 GPU): full parity suite, smoke, default bench + reference arm, cfg4 / cfg2 lines, launch list.
mkdir -p gpurun_out; S=gpurun_out/r65_summary.txt; : > $S
timeout 1700 python -m pytest tests -m gpu -q --timeout 300 > gpurun_out/r65_tests.log 2>&1; echo "tests exit $?" >> $S
tail -6 gpurun_out/r65_tests.log >> $S
grep -E "^FAILED|^ERROR" gpurun_out/r65_tests.log | head -20 >> $S
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r65_smoke.log 2>&1; echo "smoke exit $?" >> $S
timeout 600 python bench.py > gpurun_out/r65_bench_default.log 2>&1; echo "bench default exit $?" >> $S
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r65_bench_reference.log 2>&1; echo "bench reference exit $?" >> $S
timeout 600 python bench.py --config cfg4 --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/r65_bench_cfg4.log 2>&1; echo "bench cfg4 exit $?" >> $S
timeout 600 python bench.py --config cfg2 --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/r65_bench_cfg2.log 2>&1; echo "bench cfg2 exit $?" >> $S
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --profile-window"
timeout 600 ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none -c 900 --csv --log-file gpurun_out/r65_launches.csv $CMD > gpurun_out/r65_ncu_launch.log 2>&1; echo "ncu cfg3 launches exit $?" >> $S
grep -h '^{' gpurun_out/r65_bench_default.log gpurun_out/r65_bench_reference.log gpurun_out/r65_bench_cfg4.log gpurun_out/r65_bench_cfg2.log | cut -c1-260 >> $S
cat $S
