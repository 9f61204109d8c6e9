#!/bin/bash
# cluster-size sweep of the MMQ kernel on a few shapes
for cl in 1 2 4 8; do
  for a in "q8_0 14336 4096 512" "q4_K 14336 4096 512" "q8_0 128256 4096 512" "q8_0 14336 4096 2048"; do
    echo -n "cluster=$cl "; B200Q_MMQ_CLUSTER=$cl timeout 60 python tools/mmq_one.py $a 20 2>&1 | tail -1
  done
done
