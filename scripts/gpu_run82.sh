#!/bin/bash
# GPU session 82 (1 GPU): shipped NT rule of the grouped GEMM: sweep with defaults + the MoE / decode-step parity tests.
mkdir -p gpurun_out; S=gpurun_out/r82_summary.txt; : > $S
timeout 300 python benchmarks/bench_moe.py --tps 1,8 --batches 128,512,2048 > gpurun_out/r82_moe.jsonl 2>/dev/null; echo "bench_moe exit $?" >> $S
python - <<'PY' >> $S
import json
for l in open("gpurun_out/r82_moe.jsonl"):
    d = json.loads(l)
    print(d["gemm"], "tp", d["tp"], "batch", d["batch"], "ragged", d["ragged_us"], "aligned", d["aligned_us"], "ref", d.get("ref_cutlass_us"), "x", d.get("speedup_vs_ref"))
PY
timeout 400 python -m pytest tests/test_gpu_fused_moe.py tests/test_gpu_cutlass_moe.py tests/test_moe_layer.py tests/test_gpu_moe_routing.py tests/test_gpu_decode_step.py tests/test_c_abi.py -m gpu -q --timeout 120 > gpurun_out/r82_tests.log 2>&1; echo "tests exit $?" >> $S
tail -2 gpurun_out/r82_tests.log >> $S
cat $S
