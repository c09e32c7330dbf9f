#!/bin/bash
# GPU session 78 (2 GPUs): multi-GPU parity at HEAD (peer all-reduce forms, stress, TP-2 == TP-1 incl. the Mixtral path with
# the fused router / quantise-while-gathering pipeline), TP-2 bench lines of config 3 and config 4, reference arm under torchrun.
mkdir -p gpurun_out; S=gpurun_out/r78_summary.txt; : > $S
timeout 900 python -m pytest tests/test_gpu_peer_allreduce.py -m gpu -q --timeout 600 > gpurun_out/r78_tests.log 2>&1; echo "multi-gpu tests exit $?" >> $S
tail -4 gpurun_out/r78_tests.log >> $S
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 600 $TR --master-port 29542 bench.py --gpus 2 --steps 20 --warmup 3 > gpurun_out/r78_bench_tp2.log 2>&1; echo "bench tp2 exit $?" >> $S
timeout 600 $TR --master-port 29544 bench.py --gpus 2 --steps 10 --warmup 3 --config cfg4 --no-cpu-baseline > gpurun_out/r78_bench_cfg4_tp2.log 2>&1; echo "bench cfg4 tp2 exit $?" >> $S
timeout 600 $TR --master-port 29545 bench.py --impl reference --gpus 2 --steps 2 --warmup 1 > gpurun_out/r78_bench_ref_tp2.log 2>&1; echo "bench reference tp2 exit $?" >> $S
grep -h '^{' gpurun_out/r78_bench_tp2.log gpurun_out/r78_bench_cfg4_tp2.log gpurun_out/r78_bench_ref_tp2.log | cut -c1-300 >> $S
cat $S
