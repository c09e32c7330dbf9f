#!/bin/bash
# GPU session 32 (round 2, 1 GPU): first run of the round-2 tree -- PDL on every launch, fused RoPE + cache write,
# cluster-split silu_mul_quant, one seeded weight set + TP-N == TP-1 on one GPU (gloo), bench with the context
# advancing on the device + GEMM rooflines, scaled_mm next to cuBLASLt AND the reference's sm100 CUTLASS kernel.
mkdir -p gpurun_out; S=gpurun_out/r32_summary.txt; : > $S
timeout 1500 python -m pytest tests -m gpu -q --timeout 300 > gpurun_out/r32_tests.log 2>&1; echo "tests exit $?" >> $S
tail -6 gpurun_out/r32_tests.log >> $S
grep -E "FAILED|ERROR" gpurun_out/r32_tests.log | head -20 >> $S
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r32_smoke.log 2>&1; echo "smoke exit $?" >> $S
timeout 600 python bench.py > gpurun_out/r32_bench_default.log 2>&1; echo "bench default exit $?" >> $S
VB200_PDL=0 timeout 600 python bench.py --no-cpu-baseline > gpurun_out/r32_bench_nopdl.log 2>&1; echo "bench nopdl exit $?" >> $S
timeout 600 python benchmarks/bench_scaled_mm.py > gpurun_out/r32_bench_mm.log 2>&1; echo "bench_mm exit $?" >> $S
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --profile-window"
timeout 600 ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none -c 900 --csv --log-file gpurun_out/r32_launches.csv $CMD > gpurun_out/r32_ncu_launch.log 2>&1; echo "ncu launches exit $?" >> $S
grep -h '^{' gpurun_out/r32_bench_default.log gpurun_out/r32_bench_nopdl.log | cut -c1-400 >> $S
cat $S
