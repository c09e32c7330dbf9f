#!/bin/bash
# GPU session 43 (2 GPUs): multi-GPU parity (peer all-reduce one-shot/two-shot, stress, TP-2 == TP-1), per-op timing, TP-2 bench lines.
mkdir -p gpurun_out; S=gpurun_out/r43_summary.txt; : > $S
timeout 1500 python -m pytest tests/test_gpu_peer_allreduce.py -m gpu -q --timeout 900 > gpurun_out/r43_tests.log 2>&1; echo "multi-gpu tests exit $?" >> $S
tail -5 gpurun_out/r43_tests.log >> $S
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 300 $TR --master-port 29541 tests/multi_gpu/check_peer_allreduce.py > gpurun_out/r43_check_peer_ar.log 2>&1; echo "check exit $?" >> $S
grep -E "us|OK|FAIL" gpurun_out/r43_check_peer_ar.log | tail -30 >> $S
timeout 600 $TR --master-port 29542 bench.py --gpus 2 --steps 20 --warmup 3 > gpurun_out/r43_bench_tp2.log 2>&1; echo "bench tp2 exit $?" >> $S
VB200_PEER_AR=0 timeout 600 $TR --master-port 29543 bench.py --gpus 2 --steps 20 --warmup 3 > gpurun_out/r43_bench_tp2_nccl.log 2>&1; echo "bench tp2 nccl exit $?" >> $S
timeout 600 $TR --master-port 29544 bench.py --gpus 2 --steps 10 --warmup 3 --config cfg4 > gpurun_out/r43_bench_cfg4_tp2.log 2>&1; echo "bench cfg4 tp2 exit $?" >> $S
grep -h '^{' gpurun_out/r43_bench_tp2.log gpurun_out/r43_bench_tp2_nccl.log gpurun_out/r43_bench_cfg4_tp2.log | cut -c1-250 >> $S
cat $S
