#!/bin/bash
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_list.py -q -x > gpurun_out/pytest_list.log 2>&1; echo "pytest list rc=$?"; tail -4 gpurun_out/pytest_list.log
timeout 200 python tools/list_bench.py 8b > gpurun_out/list_bench.log 2>&1; echo "list_bench rc=$?"; cat gpurun_out/list_bench.log
