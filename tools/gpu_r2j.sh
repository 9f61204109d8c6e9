#!/bin/bash
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_mmq.py tests/test_gpu_parity.py -q -x -k "mxfp4" > gpurun_out/pytest_mxf.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/pytest_mxf.log
timeout 120 python tools/moe_bench.py > gpurun_out/moe_bench.log 2>&1; echo "moe_bench rc=$?"; tail -6 gpurun_out/moe_bench.log
for a in "mxfp4 2880 2880 512" "mxfp4 11520 2880 512" "mxfp4 2880 2880 64" "mxfp4 4096 4096 512"; do timeout 60 python tools/mmq_one.py $a 20 2>&1 | tail -1; done
