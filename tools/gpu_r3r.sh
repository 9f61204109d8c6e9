#!/bin/bash
# word-wise tensor-core dequantizers for q5_K / q4_1 / q5_0 / q5_1: parity, Q5_K_M prompt throughput
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_mmq.py -x -q -k "q4_1 or q5_0 or q5_1 or q5_K" > gpurun_out/pytest_mmq_dq.log 2>&1; echo "mmq rc=$?"; tail -3 gpurun_out/pytest_mmq_dq.log
D=$PWD/oracle/_ref; if grep -q avx512_vnni /proc/cpuinfo && [ -x $D/avx512/llama-bench ]; then D=$D/avx512; fi
export GGML_BACKEND_PATH=$PWD/llama.cpp-gfx906_b200/lib/libggml-b200.so
{
python tools/make_synth_gguf.py --out /tmp/synth-q5km.gguf --ftype q5_k_m
timeout 600 $D/llama-bench -m /tmp/synth-q5km.gguf -ngl 99 -fa 0 -p 512 -n 128 -r 3 2>&1 | tail -4
rm -f /tmp/synth-q5km.gguf
} > gpurun_out/llama_bench_q5km.log 2>&1
cat gpurun_out/llama_bench_q5km.log
