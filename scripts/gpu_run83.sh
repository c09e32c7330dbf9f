#!/bin/bash
# GPU session 83 (1 GPU): final state: the driver's pytest command, smoke, default bench, MoE sweep (aligned path with the new block rule).
mkdir -p gpurun_out; S=gpurun_out/r83_summary.txt; : > $S
timeout 1200 python -m pytest tests/ -x -q -m gpu > gpurun_out/r83_tests.log 2>&1; echo "pytest tests/ -x -q -m gpu exit $?" >> $S
tail -3 gpurun_out/r83_tests.log >> $S
grep -E "^FAILED|^ERROR" gpurun_out/r83_tests.log | head -20 >> $S
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r83_smoke.log 2>&1; echo "smoke exit $?" >> $S
timeout 600 python bench.py --no-cpu-baseline > gpurun_out/r83_bench_default.log 2>&1; echo "bench default exit $?" >> $S
grep -h '^{' gpurun_out/r83_bench_default.log | cut -c1-260 >> $S
timeout 200 python benchmarks/bench_moe.py --tps 1 --batches 512,2048 > gpurun_out/r83_moe.jsonl 2>/dev/null; echo "bench_moe exit $?" >> $S
python - <<'PY' >> $S
import json
for l in open("gpurun_out/r83_moe.jsonl"):
    d = json.loads(l)
    print(d["gemm"], "batch", d["batch"], "ragged", d["ragged_us"], "aligned", d["aligned_us"], "ref", d.get("ref_cutlass_us"))
PY
cat $S
