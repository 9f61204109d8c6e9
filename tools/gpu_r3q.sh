#!/bin/bash
mkdir -p gpurun_out
timeout 600 python bench.py --no-llama-bench > gpurun_out/bench_q5k.json 2> gpurun_out/bench_q5k.err; echo "bench rc=$?"; cut -c1-330 gpurun_out/bench_q5k.json
timeout 900 python -m pytest tests/test_gpu_list.py -x -q > gpurun_out/pytest_list_q5k.log 2>&1; echo "list rc=$?"; tail -3 gpurun_out/pytest_list_q5k.log
timeout 600 python bench.py --no-llama-bench --steps 100 2> /dev/null | cut -c1-200
