#!/bin/bash
mkdir -p gpurun_out
timeout 60 python tools/mmq_one.py mxfp4 2880 2880 64 3 > gpurun_out/one_plain.log 2>&1; echo "plain rc=$?"
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'mxf' -c 40 --csv --log-file gpurun_out/mxf_launches.csv python tools/mmq_one.py mxfp4 2880 2880 64 3 > gpurun_out/ncu_mxf.log 2>&1; echo "ncu rc=$?"
python - <<'PY'
import csv
rows=list(csv.reader(open('gpurun_out/mxf_launches.csv')))
hi=[i for i,r in enumerate(rows) if r and r[0]=="ID"][0]
hdr=rows[hi]; kn=hdr.index("Kernel Name"); mv=hdr.index("Metric Value"); gs=hdr.index("Grid Size")
for r in rows[hi+1:hi+13]:
    print(r[kn][:50], r[gs], r[mv])
PY
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'mxf' -c 12 --csv --log-file gpurun_out/mxf_launches2.csv python tools/moe_bench.py > gpurun_out/ncu_mxf2.log 2>&1; echo "ncu rc=$?"
python - <<'PY'
import csv
rows=list(csv.reader(open('gpurun_out/mxf_launches2.csv')))
hi=[i for i,r in enumerate(rows) if r and r[0]=="ID"][0]
hdr=rows[hi]; kn=hdr.index("Kernel Name"); mv=hdr.index("Metric Value"); gs=hdr.index("Grid Size")
for r in rows[hi+1:hi+13]:
    print(r[kn][:50], r[gs], r[mv])
PY
