#!/bin/bash
# per-kernel durations of a 4 decoded tokens through llama-bench (ncu launch list, serialised; shares, not absolutes)
mkdir -p gpurun_out
D=$PWD/oracle/_ref; if grep -q avx512_vnni /proc/cpuinfo && [ -x $D/avx512/llama-bench ]; then D=$D/avx512; fi
export GGML_BACKEND_PATH=$PWD/llama.cpp-gfx906_b200/lib/libggml-b200.so
python tools/make_synth_gguf.py --arch gpt-oss --layers 8 --out /tmp/synth-q4km.gguf
GGML_B200_NO_GRAPHS=1 timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/gptoss_tg_launches.csv \
  $D/llama-bench -m /tmp/synth-q4km.gguf -ngl 99 -fa 0 -p 0 -n 3 -r 1 --no-warmup > gpurun_out/ncu_gptoss_tg.log 2>&1
echo "ncu rc=$?"; tail -3 gpurun_out/ncu_gptoss_tg.log
python - <<'PY' | tee gpurun_out/gptoss_tg_kernels.txt
import csv, re
with open('gpurun_out/gptoss_tg_launches.csv') as f:
    lines = [l for l in f if l.startswith('"')]
agg = {}; n = 0
for row in csv.DictReader(lines):
    if row.get('Metric Name') != 'gpu__time_duration.sum': continue
    short = re.sub(r'^void\s+', '', row['Kernel Name'])
    short = re.sub(r'b200q::\(anonymous namespace\)::|b200q::|\(anonymous namespace\)::', '', short)
    short = re.sub(r'\(.*$', '', short)
    v = float(row['Metric Value'].replace(',', '')); u = row['Metric Unit']
    us = v / 1000.0 if u.startswith('n') else v
    a = agg.setdefault(short, [0, 0.0]); a[0] += 1; a[1] += us; n += 1
tot = sum(a[1] for a in agg.values())
print(f"{n} launches, {tot:.0f} us serialised")
for k, (c, t) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print(f"{k[:60]:60s} {c:6d} launches {t:9.1f} us {100*t/tot:5.1f} %  {t/c:7.2f} us each")
PY
rm -f /tmp/synth-q4km.gguf
