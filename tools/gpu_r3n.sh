#!/bin/bash
# rope_store with biases: layer tests, gpt-oss decode
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_plugin.py tests/test_gpu_llama_bench.py -x -q -s -k "layer or moe_block or llama_bench or replay" 2>&1 | grep -a "fusions\|passed\|failed\|rror\|assert" > gpurun_out/pytest_ropebias.log; grep -a "biases\|passed\|failed\|rror\|assert" gpurun_out/pytest_ropebias.log | tail -12 | cut -c1-330
D=$PWD/oracle/_ref; if grep -q avx512_vnni /proc/cpuinfo && [ -x $D/avx512/llama-bench ]; then D=$D/avx512; fi
export GGML_BACKEND_PATH=$PWD/llama.cpp-gfx906_b200/lib/libggml-b200.so
{
python tools/make_synth_gguf.py --arch gpt-oss --out /tmp/synth-gptoss.gguf
GGML_B200_STATS=1 timeout 600 $D/llama-bench -m /tmp/synth-gptoss.gguf -ngl 99 -fa 0 -p 0 -n 128 -r 3 2>&1 | tail -4
rm -f /tmp/synth-gptoss.gguf
} > gpurun_out/llama_bench_ropebias.log 2>&1
cat gpurun_out/llama_bench_ropebias.log
