#!/bin/bash
# single-GPU visit: the no-alloc test alone in a fresh process, then the whole gpu suite (it runs last there), then bench
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q > gpurun_out/pytest_zz_alone.log 2>&1; echo "zz alone rc=$?"; tail -3 gpurun_out/pytest_zz_alone.log
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/pytest_gpu.log
timeout 600 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"; tail -3 gpurun_out/bench.err; cat gpurun_out/bench.json
