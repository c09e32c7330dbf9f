#!/bin/bash
# GPU session 44 (8 GPUs): peer all-reduce one-shot vs two-shot at 8 ranks (parity + per-op timing), TP-8 == TP-1, cfg3 TP-8 / TP-4 bench lines.
mkdir -p gpurun_out; S=gpurun_out/r44_summary.txt; : > $S
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
timeout 300 $TR --nproc-per-node 8 --master-port 29551 tests/multi_gpu/check_peer_allreduce.py > gpurun_out/r44_check_peer_ar.log 2>&1; echo "check exit $?" >> $S
grep -E "timing|OK|FAIL|Error" gpurun_out/r44_check_peer_ar.log | cut -c1-1500 | tail -12 >> $S
timeout 600 $TR --nproc-per-node 8 --master-port 29552 tests/multi_gpu/check_tp_parity.py > gpurun_out/r44_tp_parity.log 2>&1; echo "tp parity exit $?" >> $S
grep -E "OK|FAIL|max" gpurun_out/r44_tp_parity.log | tail -12 | cut -c1-300 >> $S
timeout 600 $TR --nproc-per-node 8 --master-port 29553 bench.py --gpus 8 --steps 20 --warmup 3 > gpurun_out/r44_bench_tp8.log 2>&1; echo "bench tp8 exit $?" >> $S
timeout 600 $TR --nproc-per-node 4 --master-port 29554 bench.py --gpus 4 --steps 20 --warmup 3 > gpurun_out/r44_bench_tp4.log 2>&1; echo "bench tp4 exit $?" >> $S
grep -h '^{' gpurun_out/r44_bench_tp8.log gpurun_out/r44_bench_tp4.log | cut -c1-250 >> $S
cat $S
