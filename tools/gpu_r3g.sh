#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_gpu_dense.py tests/test_gpu_plugin.py tests/test_gpu_llama_bench.py -x -q > gpurun_out/pytest_moefuse.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_moefuse.log
D=$PWD/oracle/_ref; if grep -q avx512_vnni /proc/cpuinfo && [ -x $D/avx512/llama-bench ]; then D=$D/avx512; fi
export GGML_BACKEND_PATH=$PWD/llama.cpp-gfx906_b200/lib/libggml-b200.so
{
python tools/make_synth_gguf.py --arch gpt-oss --out /tmp/synth-gptoss.gguf
timeout 900 $D/llama-bench -m /tmp/synth-gptoss.gguf -ngl 99 -fa 0 -p 512 -n 128 -r 3 2>&1 | tail -4
echo "=== GGML_B200_NO_FUSE=1"
GGML_B200_NO_FUSE=1 timeout 900 $D/llama-bench -m /tmp/synth-gptoss.gguf -ngl 99 -fa 0 -p 0 -n 128 -r 3 2>&1 | tail -3
rm -f /tmp/synth-gptoss.gguf
} > gpurun_out/llama_bench_gptoss.log 2>&1
cat gpurun_out/llama_bench_gptoss.log
