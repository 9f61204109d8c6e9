#!/bin/bash
# the reference's llama-bench on a synthetic gpt-oss-20B-shaped GGUF (MXFP4 experts) through the plugin
mkdir -p gpurun_out
D=$PWD/oracle/_ref; if grep -q avx512_vnni /proc/cpuinfo && [ -x $D/avx512/llama-bench ]; then D=$D/avx512; fi
export GGML_BACKEND_PATH=$PWD/llama.cpp-gfx906_b200/lib/libggml-b200.so
{
python tools/make_synth_gguf.py --arch gpt-oss --out /tmp/synth-gptoss.gguf
echo "=== B200 plugin, gpt-oss-20B shaped, -fa 0"
timeout 900 $D/llama-bench -m /tmp/synth-gptoss.gguf -ngl 99 -fa 0 -p 512 -n 128 -r 3 2>&1 | tail -8
echo "=== CPU backend, short"
GGML_BACKEND_PATH= timeout 600 $D/llama-bench -m /tmp/synth-gptoss.gguf -ngl 0 -p 64 -n 16 -r 1 2>&1 | tail -4
rm -f /tmp/synth-gptoss.gguf
} > gpurun_out/llama_bench_gptoss.log 2>&1
cat gpurun_out/llama_bench_gptoss.log
