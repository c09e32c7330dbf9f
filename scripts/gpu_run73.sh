#!/bin/bash
# GPU session 73 (1 GPU): ragged grouped GEMM, M-block size under the new launch order.
mkdir -p gpurun_out; S=gpurun_out/r73_summary.txt; : > $S
for BM in 64 128 256; do
  VB200_MOE_RAGGED_BLOCK_M=$BM timeout 400 python benchmarks/bench_moe.py --tps 1 --batches 128,512,2048 > gpurun_out/r73_moe_bm$BM.jsonl 2> gpurun_out/r73_moe_bm$BM.err; echo "bench_moe block_m $BM exit $?" >> $S
  python - <<PY >> $S
import json
for l in open("gpurun_out/r73_moe_bm$BM.jsonl"):
    d = json.loads(l)
    print(d["gemm"], d["batch"], "ragged", d["ragged_us"], "aligned", d["aligned_us"], "ref", d.get("ref_cutlass_us"))
PY
done
cat $S
