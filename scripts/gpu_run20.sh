#!/bin/bash
# GPU session 20 (2 GPUs): multi-rank exit fix check (tiny, then cfg3) -- every command under its own short timeout.
mkdir -p gpurun_out; S=gpurun_out/r20_summary.txt; : > $S
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512"
t0=$(date +%s); timeout 150 $TR bench.py --gpus 2 --config tiny --steps 5 --warmup 3 > gpurun_out/r20_bench_tiny_tp2.log 2>&1; echo "bench tiny tp2 exit $? in $(( $(date +%s) - t0 )) s" >> $S
t0=$(date +%s); timeout 240 $TR bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/r20_bench_cfg3_tp2.log 2>&1; echo "bench cfg3 tp2 exit $? in $(( $(date +%s) - t0 )) s" >> $S
grep -h '^{' gpurun_out/r20_bench_*.log | cut -c1-300 >> $S
cat $S
