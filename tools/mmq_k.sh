#!/bin/bash
# kernel-only duration of the MMQ kernel (ncu gpu__time_duration) for debug-flag variants: tools/mmq_k.sh "0 63 127" q8_0 1024 4096 512
for f in $1; do
  B200Q_MMQ_DBG=$f ncu --metrics gpu__time_duration.sum --clock-control none -k regex:mmq_kernel -c 2 --csv --log-file /tmp/l.csv python tools/mmq_one.py $2 $3 $4 $5 2 > /dev/null 2>&1
  echo "flags=$f $2 m=$3 k=$4 n=$5: $(grep mmq_kernel /tmp/l.csv | tail -1 | awk -F'","' '{print $NF}' | tr -d '"') ns"
done
