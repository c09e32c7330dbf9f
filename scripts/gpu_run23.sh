#!/bin/bash
# GPU session 23 (8 GPUs): fused peer all-reduce at 8 / 4 ranks: parity, then bench cfg3 (TP 8, TP 4), cfg5 (70B, TP 8), cfg4 (TP 8).
mkdir -p gpurun_out; S=gpurun_out/r23_summary.txt; : > $S
TR8="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29517"
TR4="python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29518"
timeout 200 $TR8 tests/multi_gpu/check_peer_allreduce.py > gpurun_out/r23_peer_check8.log 2>&1; rc=$?; echo "peer check (8 ranks) exit $rc" >> $S
grep -h "peer all-reduce check\|FAIL\|timing" gpurun_out/r23_peer_check8.log | head -12 >> $S
if [ $rc -ne 0 ]; then grep -h "Error\|Mismatch\|Greatest" gpurun_out/r23_peer_check8.log | head -8 >> $S; echo "peer AR check failed: benches skipped" >> $S; cat $S; exit 0; fi
timeout 200 $TR8 bench.py --gpus 8 --steps 10 --warmup 3 > gpurun_out/r23_bench_cfg3_tp8_peer.log 2>&1; echo "bench cfg3 tp8 exit $?" >> $S
timeout 200 $TR4 bench.py --gpus 4 --steps 10 --warmup 3 > gpurun_out/r23_bench_cfg3_tp4_peer.log 2>&1; echo "bench cfg3 tp4 exit $?" >> $S
timeout 300 $TR8 bench.py --gpus 8 --config cfg5 --steps 10 --warmup 3 > gpurun_out/r23_bench_cfg5_tp8_peer.log 2>&1; echo "bench cfg5 tp8 exit $?" >> $S
timeout 200 $TR8 bench.py --gpus 8 --config cfg4 --steps 10 --warmup 3 > gpurun_out/r23_bench_cfg4_tp8_peer.log 2>&1; echo "bench cfg4 tp8 exit $?" >> $S
grep -h '^{' gpurun_out/r23_bench_*.log | cut -c1-230 >> $S
cat $S
