#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_plugin.py -x -q -s -k "llama_layer or moe_block or ARGSORT or ADD_ID" > gpurun_out/pytest_layer.log 2>&1; echo "rc=$?"; grep -a "llama layer\|moe block\|graph replay\|passed\|failed\|Error\|error\|not supported\|assert" gpurun_out/pytest_layer.log | tail -40
