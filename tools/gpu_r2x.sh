#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_llama_bench.py -x -q -s > gpurun_out/pytest_llama_bench.log 2>&1; echo "pytest rc=$?"; grep -a "llama-bench\|passed\|failed\|Error" gpurun_out/pytest_llama_bench.log | tail -6
timeout 900 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"; tail -3 gpurun_out/bench.err
python - <<'PY'
import json
d = json.load(open('gpurun_out/bench.json'))
print(d['value'], d['ms_per_step'], d['e2e']['value'], d['clocks'])
print(json.dumps(d['aux'].get('llama_bench_8b_q4_k_m'), indent=1))
PY
