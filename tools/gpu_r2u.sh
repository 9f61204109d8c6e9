#!/bin/bash
# programmatic dependent launch in the small kernels: parity (plugin suite + dense + zz), then llama-bench tg/pp with and without
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_gpu_plugin.py tests/test_gpu_dense.py tests/test_gpu_parity.py -x -q > gpurun_out/pytest_pdl.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/pytest_pdl.log
D=$PWD/oracle/_ref; if grep -q avx512_vnni /proc/cpuinfo && [ -x $D/avx512/llama-bench ]; then D=$D/avx512; fi
export GGML_BACKEND_PATH=$PWD/llama.cpp-gfx906_b200/lib/libggml-b200.so
{
python tools/make_synth_gguf.py --out /tmp/synth-q4km.gguf --layers 32 --ftype q4_k_m
echo "=== PDL on"
timeout 600 $D/llama-bench -m /tmp/synth-q4km.gguf -ngl 99 -fa 0 -p 512 -n 128 -r 3 2>&1 | tail -4
echo "=== PDL off in the small kernels (B200Q_NO_PDL=1)"
B200Q_NO_PDL=1 timeout 600 $D/llama-bench -m /tmp/synth-q4km.gguf -ngl 99 -fa 0 -p 512 -n 128 -r 3 2>&1 | tail -4
echo "=== PDL on, graphs off"
GGML_B200_NO_GRAPHS=1 timeout 600 $D/llama-bench -m /tmp/synth-q4km.gguf -ngl 99 -fa 0 -p 0 -n 128 -r 3 2>&1 | tail -3
rm -f /tmp/synth-q4km.gguf
} > gpurun_out/llama_bench_pdl.log 2>&1
cat gpurun_out/llama_bench_pdl.log
