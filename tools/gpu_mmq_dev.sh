#!/bin/bash
# development build of the tensor-core kernel: per-chunk timeline and knock-out ablations (B200Q_MMQ_DBG bits:
# 1 skip W loads, 2 skip dequant math, 4 skip B loads, 8 skip A stores, 16 skip proxy fence, 32 skip MMA, 64 no W pipeline, 128 no B pipeline)
mkdir -p gpurun_out
export B200Q_LIB=$PWD/llama.cpp-gfx906_b200/lib_dev/libb200q.so
{
for a in "q4_K 14336 4096 32" "q8_0 4096 4096 512" "q8_0 14336 4096 512"; do
  echo "=== timeline $a"; timeout 120 python tools/mmq_ts.py $a 2>&1 | tail -36
  for f in 0 2 8 16 24 26 32 64 66 90 128; do
    echo -n "dbg=$f "; B200Q_MMQ_DBG=$f timeout 120 python tools/mmq_one.py $a 20 2>&1 | tail -1
  done
done
} > gpurun_out/mmq_dev.log 2>&1
tail -80 gpurun_out/mmq_dev.log
