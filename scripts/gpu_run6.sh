#!/bin/bash
mkdir -p gpurun_out; S=gpurun_out/r6_summary.txt; : > $S
CMD="python benchmarks/prof_scaled_mm.py o,gate_up 256"
timeout 300 $CMD > gpurun_out/r6_plain.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:scaled_mm_kernel -s 4 -c 2 -o gpurun_out/r6_mm_o $CMD > gpurun_out/r6_ncu1.log 2>&1
echo "ncu o exit $?" >> $S
timeout 900 ncu --set full --clock-control none --import-source on -k regex:scaled_mm_kernel -s 10 -c 2 -o gpurun_out/r6_mm_gu $CMD > gpurun_out/r6_ncu2.log 2>&1
echo "ncu gu exit $?" >> $S
cat $S
