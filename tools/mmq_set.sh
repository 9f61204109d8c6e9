#!/bin/bash
# the MMQ timing set (pack + kernel per call, CUDA-graph replay): bash tools/mmq_set.sh > log
for a in "q8_0 14336 4096 512" "q8_0 4096 4096 512" "q8_0 4096 14336 512" "q8_0 128256 4096 512" "q4_K 14336 4096 512" \
         "q6_K 4096 14336 512" "q4_0 14336 4096 512" "mxfp4 11520 2880 512" "q8_0 14336 4096 2048"; do
    timeout 120 python tools/mmq_one.py $a 20 2>&1 | tail -1
done
