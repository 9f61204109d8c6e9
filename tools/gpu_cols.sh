#!/bin/bash
# GPU visit for the multi-column decode kernel: timing of planner variants
mkdir -p gpurun_out
timeout 200 python tools/quick_bench.py cols q4_K > gpurun_out/cols_a.log 2>&1; echo "cols rc=$?"
B200Q_MMA_EARLY=0 timeout 200 python tools/quick_bench.py cols q4_K > gpurun_out/cols_b.log 2>&1; echo "cols noearly rc=$?"
B200Q_MMA_KC=2 timeout 200 python tools/quick_bench.py cols q4_K > gpurun_out/cols_c.log 2>&1; echo "cols kc2 rc=$?"
B200Q_MMA_NW=8 B200Q_MMA_KC=8 timeout 200 python tools/quick_bench.py cols q4_K > gpurun_out/cols_d.log 2>&1; echo "cols nw8kc8 rc=$?"
echo "shape | default | no early prefetch | kc=2 | nw=8 kc=8"
paste <(awk '{print $1,$2,$3,$4,$5,$6,$9}' gpurun_out/cols_a.log) <(awk '{print $9}' gpurun_out/cols_b.log) <(awk '{print $9}' gpurun_out/cols_c.log) <(awk '{print $9}' gpurun_out/cols_d.log)
grep 128256 gpurun_out/cols_a.log gpurun_out/cols_b.log gpurun_out/cols_c.log gpurun_out/cols_d.log | awk '{print $1,$6,$9,$10}'
