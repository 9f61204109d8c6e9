#!/bin/bash
# llama-bench on 2 GPUs through the plugin: -sm layer (scheduler copies between the two backends) and -sm row (split buffer type)
mkdir -p gpurun_out
D=$PWD/oracle/_ref; if grep -q avx512_vnni /proc/cpuinfo && [ -x $D/avx512/llama-bench ]; then D=$D/avx512; fi
export GGML_BACKEND_PATH=$PWD/llama.cpp-gfx906_b200/lib/libggml-b200.so
timeout 600 python -m pytest tests/test_gpu_plugin.py -x -q -k "split" > gpurun_out/pytest_split.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_split.log
{
python tools/make_synth_gguf.py --out /tmp/synth-q4km.gguf --layers 32 --ftype q4_k_m
echo "=== 2 GPUs, -sm layer"
timeout 600 $D/llama-bench -m /tmp/synth-q4km.gguf -ngl 99 -fa 0 -sm layer -p 512 -n 128 -r 2 2>&1 | tail -5
echo "=== 2 GPUs, -sm row"
timeout 600 $D/llama-bench -m /tmp/synth-q4km.gguf -ngl 99 -fa 0 -sm row -p 512 -n 128 -r 2 2>&1 | tail -8
echo "=== 1 GPU (-sm none)"
timeout 600 $D/llama-bench -m /tmp/synth-q4km.gguf -ngl 99 -fa 0 -sm none -p 512 -n 128 -r 2 2>&1 | tail -4
rm -f /tmp/synth-q4km.gguf
} > gpurun_out/llama_bench_n2.log 2>&1
cat gpurun_out/llama_bench_n2.log
