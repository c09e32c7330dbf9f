#!/bin/bash
# GPU session 50 (8 GPUs): split all-reduce at 8 and 4 ranks: parity + per-op timing, cfg3 TP-8 / TP-4 bench lines (auto = split), cfg5 TP-8.
mkdir -p gpurun_out; S=gpurun_out/r50_summary.txt; : > $S
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
timeout 400 $TR --nproc-per-node 8 --master-port 29571 tests/multi_gpu/check_peer_allreduce.py > gpurun_out/r50_check_peer_ar.log 2>&1; echo "check8 exit $?" >> $S
grep -E "timing|FAIL|Error" gpurun_out/r50_check_peer_ar.log | cut -c1-1800 | tail -6 >> $S
grep -c "check: OK" gpurun_out/r50_check_peer_ar.log >> $S
timeout 400 $TR --nproc-per-node 4 --master-port 29572 tests/multi_gpu/check_peer_allreduce.py > gpurun_out/r50_check_peer_ar4.log 2>&1; echo "check4 exit $?" >> $S
grep -E "timing|FAIL|Error" gpurun_out/r50_check_peer_ar4.log | cut -c1-1800 | tail -6 >> $S
timeout 600 $TR --nproc-per-node 8 --master-port 29573 bench.py --gpus 8 --steps 20 --warmup 3 > gpurun_out/r50_bench_tp8.log 2>&1; echo "bench tp8 exit $?" >> $S
timeout 600 $TR --nproc-per-node 4 --master-port 29574 bench.py --gpus 4 --steps 20 --warmup 3 > gpurun_out/r50_bench_tp4.log 2>&1; echo "bench tp4 exit $?" >> $S
timeout 600 $TR --nproc-per-node 8 --master-port 29575 bench.py --gpus 8 --steps 10 --warmup 3 --config cfg5 > gpurun_out/r50_bench_cfg5_tp8.log 2>&1; echo "bench cfg5 tp8 exit $?" >> $S
grep -h '^{' gpurun_out/r50_bench_tp8.log gpurun_out/r50_bench_tp4.log gpurun_out/r50_bench_cfg5_tp8.log | cut -c1-250 >> $S
cat $S
