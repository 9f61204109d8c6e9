#!/bin/bash
# deferred commits in the tensor-core kernel's issue loop: parity, timings (before: see profiles/r02_mmq_accsplit_v1.txt), timeline
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_mmq.py -x -q > gpurun_out/pytest_mmq.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_mmq.log
{
for a in "q4_K 14336 4096 32" "q8_0 4096 4096 32" "q4_K 4096 4096 17" "q4_0 4096 4096 64" "q8_0 4096 4096 128" "q8_0 4096 4096 512" "q8_0 14336 4096 512" "q8_0 4096 14336 512" "q8_0 1024 4096 512" "q4_K 4096 4096 512" "q4_K 14336 4096 512" "q8_0 14336 4096 2048" "q8_0 128256 4096 512" "mxfp4 11520 2880 512"; do
  timeout 120 python tools/mmq_one.py $a 20 2>&1 | tail -1
done
B200Q_NO_MXF=1 timeout 200 python tools/moe_bench.py 2>&1 | tail -2
export B200Q_LIB=$PWD/llama.cpp-gfx906_b200/lib_dev/libb200q.so
for a in "q4_K 14336 4096 32" "q8_0 14336 4096 512"; do
  echo "=== timeline (MMA issue stamps) $a"; B200Q_MMQ_DBG=1024 timeout 120 python tools/mmq_ts.py $a 2>&1 | tail -22 | cut -c1-130
done
} > gpurun_out/mmq_defer.log 2>&1
cat gpurun_out/mmq_defer.log
