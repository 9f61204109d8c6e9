#!/bin/bash
# ncu --set full of ONE launch of the headline kernel on the whole decode step (225 matmuls, 4.63 GB): DRAM traffic per launch
mkdir -p gpurun_out
timeout 200 python tools/list_one.py 0 32 1 > gpurun_out/list_one_full.log 2>&1; echo "plain rc=$?"; cat gpurun_out/list_one_full.log
timeout 900 ncu --set full --clock-control none --import-source on -k regex:gemv_list -s 3 -c 1 -f -o gpurun_out/list_fullstep python tools/list_one.py 0 32 1 > gpurun_out/ncu_list_fullstep.log 2>&1; echo "ncu rc=$?"; tail -2 gpurun_out/ncu_list_fullstep.log
ncu -i gpurun_out/list_fullstep.ncu-rep --page raw --csv > gpurun_out/ncu_list_fullstep.csv 2>/dev/null
python - <<'PY'
import csv
rows = list(csv.reader(open('gpurun_out/ncu_list_fullstep.csv')))
hdr, units, vals = rows[0], rows[1], rows[2]
for want in ('dram__bytes_read.sum', 'dram__bytes_write.sum', 'gpu__time_duration.sum', 'dram__throughput.avg.pct_of_peak_sustained_elapsed', 'sm__throughput.avg.pct_of_peak_sustained_elapsed'):
    if want in hdr:
        i = hdr.index(want); print(want, vals[i], units[i])
PY
ls -la gpurun_out/list_fullstep.ncu-rep
