#!/bin/bash
mkdir -p gpurun_out
timeout 300 ncu --set full --clock-control none --import-source on -k regex:gemv_mma -s 4 -c 1 -f -o gpurun_out/mma_head \
    python tools/quick_bench.py one q4_K 128256 4096 8 > gpurun_out/ncu_mma_head.log 2>&1; echo "ncu head rc=$?"
timeout 300 ncu --set full --clock-control none --import-source on -k regex:gemv_mma -s 4 -c 1 -f -o gpurun_out/mma_ffn \
    python tools/quick_bench.py one q4_K 14336 4096 8 > gpurun_out/ncu_mma_ffn.log 2>&1; echo "ncu ffn rc=$?"
