#!/bin/bash
mkdir -p gpurun_out
timeout 200 python -m pytest tests/test_gpu_list.py tests/test_gpu_parity.py -q -x > gpurun_out/pytest_lp.log 2>&1; rc=$?; echo "pytest rc=$rc"; tail -3 gpurun_out/pytest_lp.log
if [ $rc -ne 0 ]; then exit 1; fi
timeout 100 python tools/list_bench.py 8b > gpurun_out/list_bench.log 2>&1; echo "list_bench rc=$?"; cat gpurun_out/list_bench.log
