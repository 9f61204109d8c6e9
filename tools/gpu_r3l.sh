#!/bin/bash
# scratch head reserved for every graph (hoisting also in MoE graphs), ADD reads the scratch: plugin tests, gpt-oss + llama decode
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_plugin.py tests/test_gpu_llama_bench.py -x -q -s 2>&1 | grep -a "fusions\|passed\|failed\|rror\|assert" > gpurun_out/pytest_scratch_all.log; grep -a "n_tokens=1 \|moe\|passed\|failed\|rror\|assert" gpurun_out/pytest_scratch_all.log | tail -12 | cut -c1-330
D=$PWD/oracle/_ref; if grep -q avx512_vnni /proc/cpuinfo && [ -x $D/avx512/llama-bench ]; then D=$D/avx512; fi
export GGML_BACKEND_PATH=$PWD/llama.cpp-gfx906_b200/lib/libggml-b200.so
{
python tools/make_synth_gguf.py --arch gpt-oss --out /tmp/synth-gptoss.gguf
timeout 600 $D/llama-bench -m /tmp/synth-gptoss.gguf -ngl 99 -fa 0 -p 512 -n 128 -r 3 2>&1 | tail -4
rm -f /tmp/synth-gptoss.gguf
python tools/make_synth_gguf.py --out /tmp/synth-q4km.gguf --ftype q4_k_m
timeout 600 $D/llama-bench -m /tmp/synth-q4km.gguf -ngl 99 -fa 0 -p 512 -n 128 -r 3 2>&1 | tail -4
rm -f /tmp/synth-q4km.gguf
} > gpurun_out/llama_bench_scratch_all.log 2>&1
cat gpurun_out/llama_bench_scratch_all.log
