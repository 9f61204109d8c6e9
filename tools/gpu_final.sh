#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/pytest_gpu.log
for c in 4 6 8; do
  B200Q_BENCH_E2E_CHUNKS=$c timeout 300 python bench.py --steps 100 --no-cpu-baseline > gpurun_out/bench_c$c.json 2>/dev/null
  python -c "import json;b=json.load(open('gpurun_out/bench_c$c.json'));print($c, b['value'],b['e2e']['value'],b['e2e']['ms_per_step'])"
done
