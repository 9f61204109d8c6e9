#!/bin/bash
mkdir -p gpurun_out
for cfg in "8 4" "8 2" "8 1" "16 1"; do
  set -- $cfg
  echo "== NW=$1 KC=$2"
  B200Q_MMA_NW=$1 B200Q_MMA_KC=$2 timeout 300 python -m pytest tests/test_gpu_parity.py -m gpu -q -k "(mma_columns) and (q6_K)" 2>&1 | grep -E "passed|failed|^FAILED" | cut -c1-110
done
