#!/bin/bash
# plugin visit: the reference's test-backend-ops against the plugin + in-process driver tests
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_plugin.py -q -x > gpurun_out/pytest_plugin.log 2>&1; echo "pytest plugin rc=$?"; tail -8 gpurun_out/pytest_plugin.log
for op in GET_ROWS MUL_MAT MUL_MAT_ID; do echo "$op: OK $(grep -c 'OK$' gpurun_out/tbo_$op.txt) unsupported $(grep -c 'not supported' gpurun_out/tbo_$op.txt) FAIL $(grep -c FAIL gpurun_out/tbo_$op.txt)"; done
grep FAIL gpurun_out/tbo_*.txt | head -20
