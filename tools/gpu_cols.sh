#!/bin/bash
# GPU visit for the multi-column decode kernel: parity first (bounded), then timing of new vs old path
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "mma or small or shapes or edge" > gpurun_out/pytest_mma.log 2>&1; echo "pytest mma rc=$?" | tee -a gpurun_out/pytest_mma.log
tail -5 gpurun_out/pytest_mma.log
timeout 200 python tools/quick_bench.py cols > gpurun_out/cols_new.log 2>&1; echo "cols rc=$?"
B200Q_MMA_KC=8 timeout 200 python tools/quick_bench.py cols q4_K > gpurun_out/cols_kc8.log 2>&1; echo "cols kc8 rc=$?"
B200Q_MMA_KC=2 timeout 200 python tools/quick_bench.py cols q4_K > gpurun_out/cols_kc2.log 2>&1; echo "cols kc2 rc=$?"
B200Q_NO_MMA_GEMV=1 timeout 200 python tools/quick_bench.py cols q4_K > gpurun_out/cols_old.log 2>&1; echo "cols old rc=$?"
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
tail -3 gpurun_out/pytest_gpu.log
cat gpurun_out/cols_new.log
