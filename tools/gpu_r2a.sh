#!/bin/bash
# round 2, first GPU visit: the persistent list kernel (parity + timing), the plugin under pytest, a short bench
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_list.py -x -q > gpurun_out/pytest_list.log 2>&1; echo "pytest list rc=$?"; tail -5 gpurun_out/pytest_list.log
timeout 200 python tools/list_bench.py 8b > gpurun_out/list_bench.log 2>&1; echo "list_bench rc=$?"; cat gpurun_out/list_bench.log
timeout 900 python -m pytest tests/test_gpu_plugin.py -x -q > gpurun_out/pytest_plugin.log 2>&1; echo "pytest plugin rc=$?"; tail -5 gpurun_out/pytest_plugin.log
timeout 400 python bench.py --steps 50 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"; tail -3 gpurun_out/bench.err; cat gpurun_out/bench.json
timeout 100 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/smoke.log
