#!/bin/bash
# GPU visit for the multi-column decode kernel: parity first (bounded), then timing with / without CTA-pair clusters
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "mma or small or shapes or edge" > gpurun_out/pytest_mma.log 2>&1; echo "pytest mma rc=$?" | tee -a gpurun_out/pytest_mma.log
tail -4 gpurun_out/pytest_mma.log
timeout 200 python tools/quick_bench.py cols > gpurun_out/cols_new.log 2>&1; echo "cols rc=$?"
B200Q_MMA_CLUSTER=1 timeout 200 python tools/quick_bench.py cols > gpurun_out/cols_nocluster.log 2>&1; echo "cols nocluster rc=$?"
paste <(awk '{print $1,$2,$3,$4,$5,$6,$9}' gpurun_out/cols_new.log) <(awk '{print $9}' gpurun_out/cols_nocluster.log)
