#!/bin/bash
# GPU session 17 (2 GPUs): fixed CUTLASS-MoE tests, then the tensor-parallel bench at N=2 (cfg3, cfg4) under torchrun.
mkdir -p gpurun_out; S=gpurun_out/r17_summary.txt; : > $S
timeout 300 python -m pytest tests/test_gpu_cutlass_moe.py -m gpu -q --timeout 60 > gpurun_out/r17_new_tests.log 2>&1; echo "new tests exit $?" >> $S
tail -5 gpurun_out/r17_new_tests.log >> $S
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511"
NCCL_DEBUG=INFO timeout 500 $TR bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/r17_bench_cfg3_tp2.log 2>&1; echo "bench cfg3 tp2 exit $?" >> $S
timeout 500 $TR bench.py --gpus 2 --config cfg4 --steps 10 --warmup 3 > gpurun_out/r17_bench_cfg4_tp2.log 2>&1; echo "bench cfg4 tp2 exit $?" >> $S
timeout 300 $TR bench.py --gpus 2 --impl reference --steps 2 --warmup 1 > gpurun_out/r17_bench_ref_tp2.log 2>&1; echo "bench ref tp2 exit $?" >> $S
grep -h '^{' gpurun_out/r17_bench_*.log | cut -c1-400 >> $S
cat $S
