mkdir -p gpurun_out
nvidia-smi > gpurun_out/smi.txt 2>&1
for f in test_gpu_cache test_gpu_quant test_gpu_moe_routing test_gpu_attention test_gpu_scaled_mm; do
  timeout 600 python -m pytest tests/$f.py -m gpu -q -x --maxfail=8 --timeout 120 > gpurun_out/r1_$f.log 2>&1
  echo "$f exit $?" >> gpurun_out/r1_summary.txt
done
timeout 300 python benchmarks/bench_paged_attention.py > gpurun_out/r1_bench_pa.log 2>&1
echo "bench_pa exit $?" >> gpurun_out/r1_summary.txt
timeout 300 python benchmarks/bench_scaled_mm.py > gpurun_out/r1_bench_mm.log 2>&1
echo "bench_mm exit $?" >> gpurun_out/r1_summary.txt
cat gpurun_out/r1_summary.txt
