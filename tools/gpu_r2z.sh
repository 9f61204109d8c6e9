#!/bin/bash
# packed-activation reuse, warp-per-row soft_max, vectorised glu: parity (plugin suite incl. test-backend-ops, mmq, parity), llama-bench
mkdir -p gpurun_out
timeout 1800 python -m pytest tests/test_gpu_plugin.py tests/test_gpu_mmq.py tests/test_gpu_llama_bench.py -x -q > gpurun_out/pytest_pack.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/pytest_pack.log
D=$PWD/oracle/_ref; if grep -q avx512_vnni /proc/cpuinfo && [ -x $D/avx512/llama-bench ]; then D=$D/avx512; fi
export GGML_BACKEND_PATH=$PWD/llama.cpp-gfx906_b200/lib/libggml-b200.so
{
python tools/make_synth_gguf.py --out /tmp/synth-q4km.gguf --layers 32 --ftype q4_k_m
timeout 600 $D/llama-bench -m /tmp/synth-q4km.gguf -ngl 99 -fa 0 -p 512,2048 -n 128 -r 3 2>&1 | tail -5
rm -f /tmp/synth-q4km.gguf
python tools/make_synth_gguf.py --out /tmp/synth-q8.gguf --layers 32 --ftype q8_0
timeout 600 $D/llama-bench -m /tmp/synth-q8.gguf -ngl 99 -fa 0 -p 512 -n 128 -r 3 2>&1 | tail -4
rm -f /tmp/synth-q8.gguf
} > gpurun_out/llama_bench_v4.log 2>&1
cat gpurun_out/llama_bench_v4.log
