#!/bin/bash
# GPU session 39 (1 GPU): feed probe with 1/2/4 consumer threads (is the 170 ns per copy the engine or the consumer chain?)
mkdir -p gpurun_out; S=gpurun_out/r39_summary.txt; : > $S
P=benchmarks/bin/l2_feed_probe; O=gpurun_out/r39_l2_feed_probe.jsonl; : > $O
for cfg in "148 1 0 0 16 8 1 1" "148 1 0 0 16 8 1 2" "148 1 0 0 16 8 2 2" "148 1 0 0 16 8 4 4" "148 1 0 0 8 16 2 2" "148 1 0 0 8 16 4 4" "148 1 0 0 4 16 4 4" "148 1 0 0 2 16 4 4" "148 1 0 0 16 12 2 2" "148 1 0 0 32 6 2 2" "148 1 0 0 64 3 2 2" "148 1 0 1 16 8 2 2" "148 1 0 1 8 16 4 4" "148 1 0 1 16 12 2 2"; do
  timeout 60 $P $cfg >> $O 2>&1 || echo "probe $cfg failed $?" >> $S
done
cat $O >> $S
cat $S
