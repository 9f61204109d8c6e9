#!/bin/bash
# round-2 ncu evidence: launch list of the bench command (our kernels), full capture of the persistent list kernel (8-layer list)
mkdir -p gpurun_out
timeout 300 python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/bench_plain.json 2> gpurun_out/bench_plain.err; echo "plain rc=$?"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'gemv|mmq|get_rows|moe_|pack' -c 400 --csv --log-file gpurun_out/launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_launch.log 2>&1; echo "ncu launches rc=$?"
timeout 100 python tools/list_one.py 0 8 > gpurun_out/list_one.log 2>&1; echo "list_one rc=$?"; cat gpurun_out/list_one.log
timeout 600 ncu --set full --clock-control none --import-source on -k regex:gemv_list -s 3 -c 1 -f -o gpurun_out/list_full python tools/list_one.py 0 8 > gpurun_out/ncu_list.log 2>&1; echo "ncu list rc=$?"; tail -3 gpurun_out/ncu_list.log
ls -la gpurun_out/*.ncu-rep
bash tools/mmq_ks.sh > gpurun_out/mmq_ks.log 2>&1; cat gpurun_out/mmq_ks.log
