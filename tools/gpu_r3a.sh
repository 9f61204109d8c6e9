#!/bin/bash
# fused decode attention: parity (plugin suite), llama-bench with / without
mkdir -p gpurun_out
timeout 1800 python -m pytest tests/test_gpu_plugin.py tests/test_gpu_llama_bench.py -x -q -s > gpurun_out/pytest_attn.log 2>&1; echo "pytest rc=$?"; grep -a "llama layer\|passed\|failed\|Error" gpurun_out/pytest_attn.log | tail -20
D=$PWD/oracle/_ref; if grep -q avx512_vnni /proc/cpuinfo && [ -x $D/avx512/llama-bench ]; then D=$D/avx512; fi
export GGML_BACKEND_PATH=$PWD/llama.cpp-gfx906_b200/lib/libggml-b200.so
{
python tools/make_synth_gguf.py --out /tmp/synth-q4km.gguf --layers 32 --ftype q4_k_m
timeout 600 $D/llama-bench -m /tmp/synth-q4km.gguf -ngl 99 -fa 0 -p 512 -n 128,512 -r 3 2>&1 | tail -5
echo "=== GGML_B200_NO_SCRATCH=1"
GGML_B200_NO_SCRATCH=1 timeout 600 $D/llama-bench -m /tmp/synth-q4km.gguf -ngl 99 -fa 0 -p 0 -n 128 -r 3 2>&1 | tail -3
echo "=== GGML_B200_NO_FUSE=1"
GGML_B200_NO_FUSE=1 timeout 600 $D/llama-bench -m /tmp/synth-q4km.gguf -ngl 99 -fa 0 -p 0 -n 128 -r 3 2>&1 | tail -3
rm -f /tmp/synth-q4km.gguf
} > gpurun_out/llama_bench_v5.log 2>&1
cat gpurun_out/llama_bench_v5.log
