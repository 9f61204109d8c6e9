#!/bin/bash
mkdir -p gpurun_out
export B200Q_LIB=$PWD/llama.cpp-gfx906_b200/lib_dev/libb200q.so
{
for a in "q4_K 14336 4096 32" "q8_0 14336 4096 512"; do
  echo "=== timeline (MMA issue stamps) $a"; B200Q_MMQ_DBG=1024 timeout 120 python tools/mmq_ts.py $a 2>&1 | tail -36
done
} > gpurun_out/mmq_dev2.log 2>&1
cat gpurun_out/mmq_dev2.log | cut -c1-130
