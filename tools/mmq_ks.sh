#!/bin/bash
# (n_sub, ksplit) sweep of the tensor-core kernel on the pp512 shapes with few row tiles
for a in "q8_0 4096 4096 512" "q8_0 1024 4096 512" "q8_0 4096 14336 512" "q8_0 14336 4096 512" "q4_K 4096 4096 512" "q8_0 4096 4096 32" "q8_0 14336 4096 32" "q8_0 4096 4096 128"; do
  for cfg in "0 0" "1 1" "1 2" "1 4" "2 1" "2 2" "2 4" "2 8"; do
    set -- $cfg
    echo -n "nsub=$1 ksplit=$2 "; B200Q_MMQ_NSUB=$1 B200Q_MMQ_KSPLIT=$2 timeout 60 python tools/mmq_one.py $a 20 2>&1 | tail -1
  done
done
