#!/bin/bash
# GPU session 51 (1 GPU): compute-sanitizer memcheck / racecheck over parts of the parity suite; ncu --set full of the dominant kernels at HEAD
# (paged attention inside the cfg3 step -> roofline.traffic; flash-layout kernel; scaled_mm at the four 8B shapes).
mkdir -p gpurun_out; S=gpurun_out/r51_summary.txt; : > $S
CS=/usr/local/cuda/bin/compute-sanitizer
timeout 900 $CS --tool memcheck --error-exitcode 9 python -m pytest tests/test_gpu_cache.py tests/test_gpu_moe_routing.py tests/test_gpu_quant.py tests/test_gpu_layer_glue.py -m gpu -q -x --timeout 800 > gpurun_out/r51_memcheck_small.log 2>&1; echo "memcheck small kernels exit $?" >> $S
tail -4 gpurun_out/r51_memcheck_small.log >> $S
timeout 900 $CS --tool memcheck --error-exitcode 9 python -m pytest tests/test_gpu_flash_attention.py tests/test_gpu_attention.py -m gpu -q -x --timeout 800 -k "S7H32 or S2H6 or backend or alibi" > gpurun_out/r51_memcheck_attn.log 2>&1; echo "memcheck attention exit $?" >> $S
tail -4 gpurun_out/r51_memcheck_attn.log >> $S
timeout 900 $CS --tool memcheck --error-exitcode 9 python -m pytest tests/test_gpu_scaled_mm.py tests/test_gpu_fused_moe.py tests/test_gpu_cutlass_moe.py -m gpu -q -x --timeout 800 -k "not vs_reference and not graph" > gpurun_out/r51_memcheck_gemm.log 2>&1; echo "memcheck gemm/moe exit $?" >> $S
tail -4 gpurun_out/r51_memcheck_gemm.log >> $S
timeout 900 $CS --tool racecheck --error-exitcode 9 python -m pytest tests/test_gpu_moe_routing.py tests/test_gpu_cache.py -m gpu -q -x --timeout 800 > gpurun_out/r51_racecheck_small.log 2>&1; echo "racecheck routing/cache exit $?" >> $S
tail -4 gpurun_out/r51_racecheck_small.log >> $S
timeout 900 $CS --tool racecheck --error-exitcode 9 python -m pytest tests/test_gpu_attention.py -m gpu -q -x --timeout 800 -k "S2H6 or S3H8" > gpurun_out/r51_racecheck_attn.log 2>&1; echo "racecheck attention exit $?" >> $S
tail -4 gpurun_out/r51_racecheck_attn.log >> $S
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --profile-window"
timeout 600 ncu --profile-from-start off --set full --clock-control none --import-source on -k regex:paged_attention_tma -c 2 -o gpurun_out/r51_pa_full $CMD > gpurun_out/r51_ncu_pa.log 2>&1; echo "ncu pa exit $?" >> $S
timeout 600 ncu --profile-from-start off --set full --clock-control none --import-source on -k regex:scaled_mm -c 4 -o gpurun_out/r51_mm_full $CMD > gpurun_out/r51_ncu_mm.log 2>&1; echo "ncu mm exit $?" >> $S
timeout 600 ncu --set full --clock-control none --import-source on -k regex:paged_attention_flash --launch-skip 6 -c 1 -o gpurun_out/r51_paf_full python benchmarks/bench_paged_attention.py --cfg 3 --flash --no-ref --iters 2 > gpurun_out/r51_ncu_paf.log 2>&1; echo "ncu flash exit $?" >> $S
for f in pa mm paf; do ncu -i gpurun_out/r51_${f}_full.ncu-rep --page raw --csv > gpurun_out/r51_${f}_raw.csv 2>/dev/null; done
cat $S
