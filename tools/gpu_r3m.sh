#!/bin/bash
D=$PWD/oracle/_ref; if grep -q avx512_vnni /proc/cpuinfo && [ -x $D/avx512/llama-bench ]; then D=$D/avx512; fi
export GGML_BACKEND_PATH=$PWD/llama.cpp-gfx906_b200/lib/libggml-b200.so
python tools/make_synth_gguf.py --arch gpt-oss --layers 4 --out /tmp/synth-gptoss.gguf
GGML_B200_STATS=1 GGML_B200_NO_GRAPHS=1 B200Q_DEBUG=1 timeout 600 $D/llama-bench -m /tmp/synth-gptoss.gguf -ngl 99 -fa 0 -p 0 -n 4 -r 1 > gpurun_out/gptoss_dbg.log 2>&1
grep "fusions" gpurun_out/gptoss_dbg.log
grep -a "^\[b200q\]" gpurun_out/gptoss_dbg.log | sed 's/0x[0-9a-f]*//g' | cut -c1-90 | sort | uniq -c | sort -rn | head -30
