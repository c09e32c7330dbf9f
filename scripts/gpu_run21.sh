#!/bin/bash
# GPU session 21 (2 GPUs): NVLink peer-memory all-reduce fused with its norm: parity vs NCCL, then bench cfg3/cfg4
# at TP 2 with it on (default) and off (VB200_PEER_AR=0).  Every command under its own short timeout.
mkdir -p gpurun_out; S=gpurun_out/r21_summary.txt; : > $S
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29514"
timeout 200 $TR tests/multi_gpu/check_peer_allreduce.py > gpurun_out/r21_peer_check.log 2>&1; rc=$?; echo "peer check exit $rc" >> $S
grep -h "peer all-reduce check\|FAIL\|Error\|error" gpurun_out/r21_peer_check.log | head -20 >> $S
if [ $rc -ne 0 ]; then tail -30 gpurun_out/r21_peer_check.log >> $S; export VB200_PEER_AR=0; echo "peer AR disabled for the bench" >> $S; fi
timeout 240 $TR bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/r21_bench_cfg3_tp2_peer.log 2>&1; echo "bench cfg3 tp2 (peer AR) exit $?" >> $S
VB200_PEER_AR=0 timeout 240 $TR bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/r21_bench_cfg3_tp2_nccl.log 2>&1; echo "bench cfg3 tp2 (NCCL) exit $?" >> $S
timeout 240 $TR bench.py --gpus 2 --config cfg4 --steps 10 --warmup 3 > gpurun_out/r21_bench_cfg4_tp2_peer.log 2>&1; echo "bench cfg4 tp2 (peer AR) exit $?" >> $S
grep -h '^{' gpurun_out/r21_bench_*.log | cut -c1-260 >> $S
cat $S
