#!/bin/bash
# GPU session 79 (1 GPU): paged attention on the one-kv-head shapes (config-5 shard, TP-8 / TP-4 shards of config 3):
# ring depth, partition size and physically consecutive blocks.
mkdir -p gpurun_out; S=gpurun_out/r79_summary.txt; : > $S
run() { echo "## $1" >> $S; shift; env "$@" 2>> gpurun_out/r79.err | python -c "
import sys, json
for l in sys.stdin:
    if l.startswith('{'):
        d = json.loads(l); print(d['cfg'], d['ours_us'], d['ours_best_us'], [v for k, v in d.items() if k.startswith('frac_of')][0])" >> $S; }
B="python benchmarks/bench_paged_attention.py --no-ref --cfg 5,3tp8,3tp4 --iters 30"
run "default" $B
for ST in 4 8 16; do run "stages $ST" VB200_PA_STAGES=$ST $B; done
for P in 512 1024 1376 4096 8192; do run "partition $P" $B --partition $P; done
run "consecutive blocks" PA_SEQ_BLOCKS=1 $B
cat $S
