#!/bin/bash
# list kernel iteration: parity tests of the list, per-shape steady state, whole pass with / without dependencies
mkdir -p gpurun_out
timeout 120 python -m pytest tests/test_gpu_list.py -q -x > gpurun_out/pytest_list.log 2>&1; rc=$?; echo "pytest list rc=$rc"; tail -4 gpurun_out/pytest_list.log
if [ $rc -ne 0 ]; then exit 1; fi
timeout 100 python tools/list_bench.py 8b > gpurun_out/list_bench.log 2>&1; echo "list_bench rc=$?"; cat gpurun_out/list_bench.log
timeout 100 python tools/list_shapes.py 8 > gpurun_out/list_shapes.log 2>&1; cat gpurun_out/list_shapes.log
