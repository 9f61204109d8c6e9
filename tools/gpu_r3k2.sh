#!/bin/bash
mkdir -p gpurun_out
timeout 300 python tools/ffn_list_bench.py 2>&1 | tail -3
timeout 600 python -m pytest tests/test_gpu_list.py -x -q -k swiglu 2>&1 | tail -2
D=$PWD/oracle/_ref; if grep -q avx512_vnni /proc/cpuinfo && [ -x $D/avx512/llama-bench ]; then D=$D/avx512; fi
export GGML_BACKEND_PATH=$PWD/llama.cpp-gfx906_b200/lib/libggml-b200.so
python tools/make_synth_gguf.py --out /tmp/synth-q4km.gguf --ftype q4_k_m
timeout 600 $D/llama-bench -m /tmp/synth-q4km.gguf -ngl 99 -fa 0 -p 0 -n 128 -r 3 2>&1 | tail -3
B200Q_DEBUG=1 timeout 600 $D/llama-bench -m /tmp/synth-q4km.gguf -ngl 99 -fa 0 -p 0 -n 2 -r 1 2>&1 | grep "gemv_list" | sort | uniq -c | sort -rn | head -8
