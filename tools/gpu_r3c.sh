#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_rowsplit.py -x -q > gpurun_out/pytest_ep_n2.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/pytest_ep_n2.log
