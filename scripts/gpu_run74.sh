#!/bin/bash
# GPU session 74 (1 GPU): grouped MoE GEMM with two weight tiles per CTA (NT = 2): parity with the variant forced on
# every shape that allows it, then A/B timing.
mkdir -p gpurun_out; S=gpurun_out/r74_summary.txt; : > $S
VB200_MOE_NT=2 timeout 300 python -m pytest tests/test_gpu_fused_moe.py tests/test_gpu_cutlass_moe.py tests/test_moe_layer.py -m gpu -q --timeout 120 -x > gpurun_out/r74_tests_moe_nt2.log 2>&1; echo "moe tests NT=2 forced exit $?" >> $S
tail -5 gpurun_out/r74_tests_moe_nt2.log >> $S
timeout 300 python -m pytest tests/test_gpu_fused_moe.py tests/test_gpu_cutlass_moe.py tests/test_moe_layer.py -m gpu -q --timeout 120 > gpurun_out/r74_tests_moe.log 2>&1; echo "moe tests default exit $?" >> $S
tail -3 gpurun_out/r74_tests_moe.log >> $S
for NT in 1 auto; do
  if [ $NT = auto ]; then unset VB200_MOE_NT; else export VB200_MOE_NT=$NT; fi
  timeout 300 python benchmarks/bench_moe.py --tps 1,8 --batches 128,512,2048 > gpurun_out/r74_moe_nt$NT.jsonl 2> gpurun_out/r74_moe_nt$NT.err; echo "bench_moe NT=$NT exit $?" >> $S
  python - <<PY >> $S
import json
for l in open("gpurun_out/r74_moe_nt$NT.jsonl"):
    d = json.loads(l)
    print(d["gemm"], "tp", d["tp"], "batch", d["batch"], "ragged", d["ragged_us"], "aligned", d["aligned_us"], "ref", d.get("ref_cutlass_us"), "rel", round(d.get("max_rel_diff_vs_ref", 0), 4))
PY
done
cat $S
