#!/bin/bash
# The reference's own llama-bench (compiled unmodified: `make -C oracle llama`) on the B200 plugin, synthetic-weight GGUFs of
# Llama-3-8B's architecture (tools/make_synth_gguf.py).  usage: tools/llama_bench_b200.sh [layers]
mkdir -p gpurun_out
L=${1:-32}
D=$PWD/oracle/_ref; if grep -q avx512_vnni /proc/cpuinfo && [ -x $D/avx512/llama-bench ]; then D=$D/avx512; fi
export GGML_BACKEND_PATH=$PWD/llama.cpp-gfx906_b200/lib/libggml-b200.so
{
python tools/make_synth_gguf.py --out /tmp/synth-q4km.gguf --layers $L --ftype q4_k_m
echo "=== B200 plugin, Q4_K_M, -fa 0"
timeout 900 $D/llama-bench -m /tmp/synth-q4km.gguf -ngl 99 -fa 0 -p 512 -n 128 -r 3 2>&1 | tail -8
echo "=== B200 plugin, Q4_K_M, CUDA graphs off (GGML_B200_NO_GRAPHS=1)"
GGML_B200_NO_GRAPHS=1 timeout 900 $D/llama-bench -m /tmp/synth-q4km.gguf -ngl 99 -fa 0 -p 512 -n 128 -r 3 2>&1 | tail -5
echo "=== B200 plugin, Q4_K_M, longer prompt"
timeout 900 $D/llama-bench -m /tmp/synth-q4km.gguf -ngl 99 -fa 0 -p 2048 -n 0 -r 2 2>&1 | tail -4
echo "=== reference CPU backend (-ngl 0), Q4_K_M, short"
GGML_BACKEND_PATH= timeout 900 $D/llama-bench -m /tmp/synth-q4km.gguf -ngl 0 -p 64 -n 16 -r 1 2>&1 | tail -5
rm -f /tmp/synth-q4km.gguf
python tools/make_synth_gguf.py --out /tmp/synth-q8.gguf --layers $L --ftype q8_0
echo "=== B200 plugin, Q8_0 (BASELINE config 3), -fa 0"
timeout 900 $D/llama-bench -m /tmp/synth-q8.gguf -ngl 99 -fa 0 -p 512 -n 128 -r 3 2>&1 | tail -5
rm -f /tmp/synth-q8.gguf
} > gpurun_out/llama_bench.log 2>&1
cat gpurun_out/llama_bench.log
