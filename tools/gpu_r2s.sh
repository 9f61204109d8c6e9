#!/bin/bash
# CUDA-graph replay in the plugin: parity + timing
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_plugin.py -x -q -s -k "graph_replay or decode_run" > gpurun_out/pytest_graphs.log 2>&1; echo "rc=$?"; grep -a "graph replay\|passed\|failed\|Error\|error" gpurun_out/pytest_graphs.log | tail -12
