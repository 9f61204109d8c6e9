#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_mmq.py tests/test_gpu_plugin.py tests/test_zz_no_alloc.py -q -x > gpurun_out/pytest_mp.log 2>&1; rc=$?; echo "pytest rc=$rc"; tail -5 gpurun_out/pytest_mp.log
for a in "q8_0 4096 4096 512" "q8_0 1024 4096 512" "q8_0 4096 14336 512" "q8_0 14336 4096 512" "q4_K 4096 4096 512" "q8_0 4096 4096 32" "q8_0 14336 4096 32" "q8_0 4096 4096 128" "q4_K 14336 4096 64"; do
  timeout 60 python tools/mmq_one.py $a 20 2>&1 | tail -1
done
timeout 200 python tools/pp512_bench.py q8_0 512 2>&1 | tail -8
