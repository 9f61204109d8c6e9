#!/bin/bash
# rope + cache store fusion: layer tests (fusion counters), llama-bench decode on Llama Q4_K_M and gpt-oss
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_plugin.py tests/test_gpu_llama_bench.py -x -q -s -k "llama_layer or moe_block or llama_bench or replay" 2>&1 | grep -a "fusions\|passed\|failed\|rror\|assert" > gpurun_out/pytest_ropestore.log; tail -25 gpurun_out/pytest_ropestore.log
D=$PWD/oracle/_ref; if grep -q avx512_vnni /proc/cpuinfo && [ -x $D/avx512/llama-bench ]; then D=$D/avx512; fi
export GGML_BACKEND_PATH=$PWD/llama.cpp-gfx906_b200/lib/libggml-b200.so
{
python tools/make_synth_gguf.py --out /tmp/synth-q4km.gguf --ftype q4_k_m
timeout 600 $D/llama-bench -m /tmp/synth-q4km.gguf -ngl 99 -fa 0 -p 512 -n 128,512 -r 3 2>&1 | tail -5
rm -f /tmp/synth-q4km.gguf
python tools/make_synth_gguf.py --arch gpt-oss --out /tmp/synth-gptoss.gguf
timeout 600 $D/llama-bench -m /tmp/synth-gptoss.gguf -ngl 99 -fa 0 -p 0 -n 128 -r 3 2>&1 | tail -3
rm -f /tmp/synth-gptoss.gguf
} > gpurun_out/llama_bench_ropestore.log 2>&1
cat gpurun_out/llama_bench_ropestore.log
