#!/bin/bash
# GPU session 16: CUTLASS-style MoE API (new), full suite, MoE microbench vs the reference's sm100 CUTLASS grouped
# GEMM, Mixtral bench (cfg4), launch list of the timed window + ncu --set full of the two dominant kernels.
mkdir -p gpurun_out; S=gpurun_out/r16_summary.txt; : > $S
timeout 400 python -m pytest tests/test_gpu_cutlass_moe.py tests/test_gpu_torch_ops.py -m gpu -q --timeout 60 -x > gpurun_out/r16_new_tests.log 2>&1; echo "new tests exit $?" >> $S
tail -15 gpurun_out/r16_new_tests.log >> $S
timeout 900 python -m pytest tests -m gpu -q --timeout 120 > gpurun_out/r16_tests.log 2>&1; echo "tests exit $?" >> $S
tail -4 gpurun_out/r16_tests.log >> $S
timeout 300 python benchmarks/bench_moe.py > gpurun_out/r16_bench_moe.log 2>&1; echo "bench_moe exit $?" >> $S
timeout 400 python bench.py --config cfg4 --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/r16_bench_cfg4.log 2>&1; echo "bench cfg4 exit $?" >> $S
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --profile-window"
timeout 600 ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none -c 900 --csv --log-file gpurun_out/r16_launches.csv $CMD > gpurun_out/r16_ncu_launch.log 2>&1; echo "ncu launches exit $?" >> $S
timeout 600 ncu --profile-from-start off --set full --clock-control none --import-source on -k regex:paged_attention -c 2 -o gpurun_out/r16_pa_full $CMD > gpurun_out/r16_ncu_pa.log 2>&1; echo "ncu pa exit $?" >> $S
timeout 600 ncu --profile-from-start off --set full --clock-control none --import-source on -k regex:scaled_mm -c 4 -o gpurun_out/r16_mm_full $CMD > gpurun_out/r16_ncu_mm.log 2>&1; echo "ncu mm exit $?" >> $S
cat $S
