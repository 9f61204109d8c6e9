#!/bin/bash
# MoE router fusion: moe block tests, gpt-oss decode + prefill
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_plugin.py tests/test_gpu_llama_bench.py -x -q -s -k "moe_block or llama_bench or ARGSORT or SOFT_MAX or ADD" 2>&1 | grep -a "fusions\|passed\|failed\|rror\|assert" > gpurun_out/pytest_route.log; tail -8 gpurun_out/pytest_route.log | cut -c1-330
D=$PWD/oracle/_ref; if grep -q avx512_vnni /proc/cpuinfo && [ -x $D/avx512/llama-bench ]; then D=$D/avx512; fi
export GGML_BACKEND_PATH=$PWD/llama.cpp-gfx906_b200/lib/libggml-b200.so
{
python tools/make_synth_gguf.py --arch gpt-oss --out /tmp/synth-gptoss.gguf
GGML_B200_STATS=1 timeout 600 $D/llama-bench -m /tmp/synth-gptoss.gguf -ngl 99 -fa 0 -p 512 -n 128 -r 3 2>&1 | tail -5
rm -f /tmp/synth-gptoss.gguf
} > gpurun_out/llama_bench_route.log 2>&1
cat gpurun_out/llama_bench_route.log
