#!/bin/bash
# GPU visit for the multi-column decode kernel: parity first (bounded), then the column sweep
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "mma or small or shapes or edge" > gpurun_out/pytest_mma.log 2>&1; echo "pytest mma rc=$?" | tee -a gpurun_out/pytest_mma.log
tail -3 gpurun_out/pytest_mma.log
timeout 200 python tools/quick_bench.py cols > gpurun_out/cols_all.log 2>&1; echo "cols rc=$?"
for t in q8_0 q4_0 q5_K; do timeout 100 python tools/quick_bench.py cols $t >> gpurun_out/cols_all.log 2>&1; done
awk '{print $1,$2,$3,$4,$5,$6,$7,$8,$9,$10,$11,$12,$13,$14}' gpurun_out/cols_all.log
timeout 600 bash tools/run_tbo.sh test -o MUL_MAT > gpurun_out/tbo_MUL_MAT.txt 2>&1; echo "tbo MUL_MAT rc=$?"; tail -3 gpurun_out/tbo_MUL_MAT.txt
