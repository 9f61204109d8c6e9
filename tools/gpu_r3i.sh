#!/bin/bash
# SWIGLU + ffn_down inside the gate / up list: list tests, layer tests (fusion counters), llama-bench decode, headline bench
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_list.py -x -q > gpurun_out/pytest_list_glu.log 2>&1; echo "list rc=$?"; tail -3 gpurun_out/pytest_list_glu.log
timeout 900 python -m pytest tests/test_gpu_plugin.py tests/test_gpu_llama_bench.py -x -q -s -k "llama_layer or moe_block or llama_bench or replay" 2>&1 | grep -a "fusions\|passed\|failed\|rror\|assert" > gpurun_out/pytest_glulist.log; tail -8 gpurun_out/pytest_glulist.log | cut -c1-330
D=$PWD/oracle/_ref; if grep -q avx512_vnni /proc/cpuinfo && [ -x $D/avx512/llama-bench ]; then D=$D/avx512; fi
export GGML_BACKEND_PATH=$PWD/llama.cpp-gfx906_b200/lib/libggml-b200.so
{
python tools/make_synth_gguf.py --out /tmp/synth-q4km.gguf --ftype q4_k_m
timeout 600 $D/llama-bench -m /tmp/synth-q4km.gguf -ngl 99 -fa 0 -p 512 -n 128,512 -r 3 2>&1 | tail -5
rm -f /tmp/synth-q4km.gguf
python tools/make_synth_gguf.py --out /tmp/synth-q8.gguf --ftype q8_0
timeout 600 $D/llama-bench -m /tmp/synth-q8.gguf -ngl 99 -fa 0 -p 0 -n 128 -r 3 2>&1 | tail -3
rm -f /tmp/synth-q8.gguf
} > gpurun_out/llama_bench_glulist.log 2>&1
cat gpurun_out/llama_bench_glulist.log
timeout 600 python bench.py --no-llama-bench > gpurun_out/bench_glulist.json 2> gpurun_out/bench_glulist.err; echo "bench rc=$?"; cut -c1-400 gpurun_out/bench_glulist.json
