#!/bin/bash
# ncu --set full of the persistent list kernel on an 8-layer list (no dependencies), source-level stall samples
mkdir -p gpurun_out
timeout 600 ncu --set full --clock-control none --import-source on -k regex:gemv_list -s 3 -c 1 -f -o gpurun_out/list_full \
    python tools/list_timeline.py ${1:-0} 8 > gpurun_out/ncu_list.log 2>&1; echo "ncu list rc=$?"; tail -3 gpurun_out/ncu_list.log
