#!/bin/bash
# GPU session 31 (1 GPU): MoE GEMM with block_m capped at 128: parity + microbench at larger batches.
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_fused_moe.py tests/test_gpu_cutlass_moe.py -m gpu -q --timeout 60 2>&1 | tail -3
timeout 200 python benchmarks/bench_moe.py --batches 128,512,2048 --tps 1 > gpurun_out/r31_bench_moe.log 2>&1; echo "bench_moe exit $?"
python - <<'PY'
import json
for l in open('gpurun_out/r31_bench_moe.log'):
    if l.startswith('{'):
        d=json.loads(l); print(d['gemm'], d['batch'], 'ragged', d['ragged_us'], 'aligned', d['aligned_us'], 'ref', d.get('ref_cutlass_us'), d.get('speedup_vs_ref'))
PY
