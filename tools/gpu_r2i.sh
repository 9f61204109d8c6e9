#!/bin/bash
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_plugin.py -q -x -k "decode_run" > gpurun_out/pytest_dr.log 2>&1; echo "pytest rc=$?"; grep -n "AssertionError\|passed\|failed" gpurun_out/pytest_dr.log | head
timeout 120 python tools/mxf_probe.py 128 > gpurun_out/mxf_probe_128.log 2>&1; echo "probe128 rc=$?"; cat gpurun_out/mxf_probe_128.log | tail -8
timeout 120 python tools/mxf_probe.py 256 > gpurun_out/mxf_probe_256.log 2>&1; echo "probe256 rc=$?"; cat gpurun_out/mxf_probe_256.log | tail -8
