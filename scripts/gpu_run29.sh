#!/bin/bash
# GPU session 29 (1 GPU): config-5 attention partition sweep with back-to-back timing.
mkdir -p gpurun_out; : > gpurun_out/r29_pa_sweep.log
for P in 0 512 704 928 1056 1376 1664 2048 2752; do
  echo "partition $P" >> gpurun_out/r29_pa_sweep.log
  timeout 60 python benchmarks/bench_paged_attention.py --cfg 5 --no-ref --partition $P >> gpurun_out/r29_pa_sweep.log 2>&1
done
python - <<'PY'
import json
cur=None
for l in open('gpurun_out/r29_pa_sweep.log'):
    if l.startswith('partition'): cur=l.split()[1]
    elif l.startswith('{'):
        d=json.loads(l); print(cur, d['ours_us'], d['ours_best_us'], d['frac_of_measured_peak'])
PY
