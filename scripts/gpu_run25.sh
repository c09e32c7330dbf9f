#!/bin/bash
# GPU session 25 (8 GPUs): fused peer all-reduce with direct loads (new default): parity at 8 ranks, bench cfg3 / cfg4 at TP 8.
mkdir -p gpurun_out; S=gpurun_out/r25_summary.txt; : > $S
TR8="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29519"
PEER_AR_TIMING=0 timeout 150 $TR8 tests/multi_gpu/check_peer_allreduce.py > gpurun_out/r25_peer_check8.log 2>&1; rc=$?; echo "peer check (8 ranks) exit $rc" >> $S
timeout 150 $TR8 bench.py --gpus 8 --steps 10 --warmup 3 > gpurun_out/r25_bench_cfg3_tp8_peer.log 2>&1; echo "bench cfg3 tp8 exit $?" >> $S
timeout 150 $TR8 bench.py --gpus 8 --config cfg4 --steps 10 --warmup 3 > gpurun_out/r25_bench_cfg4_tp8_peer.log 2>&1; echo "bench cfg4 tp8 exit $?" >> $S
grep -h '^{' gpurun_out/r25_bench_*.log | cut -c1-230 >> $S
cat $S
