#!/bin/bash
# GPU session 81 (1 GPU): NT = 2 grouped GEMM with two CTAs per SM (two-stage ring) on the short-K (w13) shapes.
mkdir -p gpurun_out; S=gpurun_out/r81_summary.txt; : > $S
show() { python - "$1" <<'PY' >> gpurun_out/r81_summary.txt
import json, sys
for l in open(sys.argv[1]):
    d = json.loads(l)
    print(d["gemm"], "tp", d["tp"], "batch", d["batch"], "ragged", d["ragged_us"], "aligned", d["aligned_us"], "ref", d.get("ref_cutlass_us"))
PY
}
echo "## NT=1" >> $S; VB200_MOE_NT=1 timeout 300 python benchmarks/bench_moe.py --tps 1,8 --batches 512,2048 > gpurun_out/r81_nt1.jsonl 2>/dev/null; show gpurun_out/r81_nt1.jsonl
echo "## NT=2, two CTAs per SM where block_m <= 128" >> $S; VB200_MOE_NT=2 timeout 300 python benchmarks/bench_moe.py --tps 1,8 --batches 512,2048 > gpurun_out/r81_nt2_2cta.jsonl 2>/dev/null; show gpurun_out/r81_nt2_2cta.jsonl
echo "## NT=2, one CTA per SM" >> $S; VB200_MOE_NT=2 VB200_MOE_NT2_2CTA=0 timeout 300 python benchmarks/bench_moe.py --tps 1,8 --batches 512,2048 > gpurun_out/r81_nt2_1cta.jsonl 2>/dev/null; show gpurun_out/r81_nt2_1cta.jsonl
VB200_MOE_NT=2 timeout 300 python -m pytest tests/test_gpu_fused_moe.py tests/test_gpu_cutlass_moe.py -m gpu -q --timeout 120 -x > gpurun_out/r81_tests_nt2.log 2>&1; echo "moe tests NT=2 forced exit $?" >> $S
tail -2 gpurun_out/r81_tests_nt2.log >> $S
cat $S
