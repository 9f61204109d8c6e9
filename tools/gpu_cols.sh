#!/bin/bash
# GPU visit for the multi-column decode kernel: parity first (bounded), then timing, then one ncu capture
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
tail -4 gpurun_out/pytest_gpu.log
timeout 200 python tools/quick_bench.py cols > gpurun_out/cols_new.log 2>&1; echo "cols rc=$?"
timeout 200 python tools/quick_bench.py cols q8_0 > gpurun_out/cols_q8_0.log 2>&1; echo "cols q8_0 rc=$?"
timeout 200 python tools/quick_bench.py cols q4_0 > gpurun_out/cols_q4_0.log 2>&1; echo "cols q4_0 rc=$?"
timeout 200 python tools/quick_bench.py cols q5_K > gpurun_out/cols_q5_K.log 2>&1; echo "cols q5_K rc=$?"
timeout 300 ncu --set full --clock-control none --import-source on -k regex:gemv_mma -s 4 -c 1 -f -o gpurun_out/mma_head \
    python tools/quick_bench.py one q4_K 128256 4096 8 > gpurun_out/ncu_mma_head.log 2>&1; echo "ncu head rc=$?"
timeout 300 ncu --set full --clock-control none --import-source on -k regex:gemv_mma -s 4 -c 1 -f -o gpurun_out/mma_ffn \
    python tools/quick_bench.py one q4_K 14336 4096 8 > gpurun_out/ncu_mma_ffn.log 2>&1; echo "ncu ffn rc=$?"
cat gpurun_out/cols_new.log
