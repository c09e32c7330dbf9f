#!/bin/bash
# GPU session 49 (2 GPUs): split all-reduce (reduce-scatter + consumer | all-gather): parity vs one-shot (bit-exact), stress, TP-2 == TP-1, timing.
mkdir -p gpurun_out; S=gpurun_out/r49_summary.txt; : > $S
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 400 $TR --master-port 29561 tests/multi_gpu/check_peer_allreduce.py > gpurun_out/r49_check_peer_ar.log 2>&1; echo "check exit $?" >> $S
grep -E "timing|OK|FAIL|Error|error" gpurun_out/r49_check_peer_ar.log | cut -c1-1800 | tail -12 >> $S
timeout 600 $TR --master-port 29562 tests/multi_gpu/stress_peer_allreduce.py > gpurun_out/r49_stress.log 2>&1; echo "stress exit $?" >> $S
tail -3 gpurun_out/r49_stress.log | cut -c1-300 >> $S
timeout 600 $TR --master-port 29563 tests/multi_gpu/check_tp_parity.py > gpurun_out/r49_tp_parity.log 2>&1; echo "tp parity exit $?" >> $S
grep -E "split|OK|FAIL" gpurun_out/r49_tp_parity.log | tail -8 | cut -c1-300 >> $S
VB200_PEER_AR_ALGO=split timeout 600 $TR --master-port 29564 bench.py --gpus 2 --steps 20 --warmup 3 > gpurun_out/r49_bench_tp2_split.log 2>&1; echo "bench tp2 split exit $?" >> $S
grep -h '^{' gpurun_out/r49_bench_tp2_split.log | cut -c1-250 >> $S
cat $S
