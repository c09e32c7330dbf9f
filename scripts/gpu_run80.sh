#!/bin/bash
# GPU session 80 (1 GPU): round-end state at HEAD with the driver's own commands: pytest -x -q -m gpu, smoke, both bench
# arms, ncu launch list of the timed window.
mkdir -p gpurun_out; S=gpurun_out/r80_summary.txt; : > $S
timeout 1500 python -m pytest tests/ -x -q -m gpu > gpurun_out/r80_tests.log 2>&1; echo "pytest tests/ -x -q -m gpu exit $?" >> $S
tail -5 gpurun_out/r80_tests.log >> $S
grep -E "^FAILED|^ERROR" gpurun_out/r80_tests.log | head -20 >> $S
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r80_smoke.log 2>&1; echo "smoke exit $?" >> $S
tail -1 gpurun_out/r80_smoke.log >> $S
timeout 600 python bench.py --impl reference --gpus 1 --steps 3 --warmup 1 > gpurun_out/r80_bench_reference.log 2>&1; echo "bench reference exit $?" >> $S
timeout 600 python bench.py > gpurun_out/r80_bench_default.log 2>&1; echo "bench default exit $?" >> $S
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --profile-window"
timeout 600 ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none -c 900 --csv --log-file gpurun_out/r80_launches.csv $CMD > gpurun_out/r80_ncu_launch.log 2>&1; echo "ncu cfg3 launches exit $?" >> $S
python scripts/agg_launches.py gpurun_out/r80_launches.csv | head -14 | cut -c1-140 >> $S
grep -h '^{' gpurun_out/r80_bench_default.log gpurun_out/r80_bench_reference.log | cut -c1-330 >> $S
cat $S
