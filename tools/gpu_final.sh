#!/bin/bash
mkdir -p gpurun_out
for cfg in "2 16" "2 32" "2 8" "99 8"; do
  set -- $cfg
  B200Q_BENCH_E2E_JOINS=$1 B200Q_BENCH_E2E_CHUNKS=$2 timeout 300 python bench.py --steps 100 --no-cpu-baseline > gpurun_out/bench_j$1c$2.json 2> gpurun_out/bench_j$1c$2.err
  python -c "import json;b=json.load(open('gpurun_out/bench_j$1c$2.json'));print('joins',$1,'chunks',$2, b['value'],b['e2e']['value'],b['e2e']['ms_per_step'],b['e2e']['mode'][:12])"
done
