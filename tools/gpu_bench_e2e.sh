#!/bin/bash
mkdir -p gpurun_out
timeout 300 python bench.py --steps 100 --no-cpu-baseline > gpurun_out/bench_p.json 2> gpurun_out/bench_p.err; echo rc=$?; tail -3 gpurun_out/bench_p.err
python -c "import json;b=json.load(open('gpurun_out/bench_p.json'));print(b['value'],b['ms_per_step'],b['e2e'])"
B200Q_BENCH_E2E=serial timeout 300 python bench.py --steps 100 --no-cpu-baseline > gpurun_out/bench_s.json 2>/dev/null
python -c "import json;b=json.load(open('gpurun_out/bench_s.json'));print(b['value'],b['e2e']['value'],b['e2e']['ms_per_step'],b['e2e']['mode'])"
