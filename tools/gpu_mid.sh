#!/bin/bash
mkdir -p gpurun_out
for t in q4_K q8_0; do
  B200Q_MMQ_MIN_N=9 timeout 300 python tools/quick_bench.py mid $t > gpurun_out/mid_${t}_mmq.log 2>&1; echo "mmq $t rc=$?"
  B200Q_MMQ_MIN_N=100000 timeout 300 python tools/quick_bench.py mid $t > gpurun_out/mid_${t}_gemv.log 2>&1; echo "gemv $t rc=$?"
done
paste <(awk '{print $1,$2,$3,$4,$5,$6,$9}' gpurun_out/mid_q4_K_mmq.log) <(awk '{print $9}' gpurun_out/mid_q4_K_gemv.log)
paste <(awk '{print $1,$2,$3,$4,$5,$6,$9}' gpurun_out/mid_q8_0_mmq.log) <(awk '{print $9}' gpurun_out/mid_q8_0_gemv.log)
