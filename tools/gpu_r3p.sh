#!/bin/bash
# q5_K in the list kernel: list tests, headline bench, llama-bench on a Q5_K_M-shaped file
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_list.py -x -q > gpurun_out/pytest_list_q5k.log 2>&1; echo "list rc=$?"; tail -3 gpurun_out/pytest_list_q5k.log
timeout 600 python bench.py --no-llama-bench > gpurun_out/bench_q5k.json 2> gpurun_out/bench_q5k.err; echo "bench rc=$?"; cut -c1-330 gpurun_out/bench_q5k.json
D=$PWD/oracle/_ref; if grep -q avx512_vnni /proc/cpuinfo && [ -x $D/avx512/llama-bench ]; then D=$D/avx512; fi
export GGML_BACKEND_PATH=$PWD/llama.cpp-gfx906_b200/lib/libggml-b200.so
{
python tools/make_synth_gguf.py --out /tmp/synth-q5km.gguf --ftype q5_k_m
GGML_B200_STATS=1 timeout 600 $D/llama-bench -m /tmp/synth-q5km.gguf -ngl 99 -fa 0 -p 512 -n 128 -r 3 2>&1 | tail -5
echo "=== GGML_B200_NO_LIST=1"
GGML_B200_NO_LIST=1 timeout 600 $D/llama-bench -m /tmp/synth-q5km.gguf -ngl 99 -fa 0 -p 0 -n 128 -r 3 2>&1 | tail -3
rm -f /tmp/synth-q5km.gguf
} > gpurun_out/llama_bench_q5km.log 2>&1
cat gpurun_out/llama_bench_q5km.log
