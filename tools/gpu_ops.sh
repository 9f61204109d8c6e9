#!/bin/bash
# the reference's test-backend-ops on the operators around the matmuls
mkdir -p gpurun_out
D=oracle/_ref; if grep -q avx512_vnni /proc/cpuinfo && [ -x oracle/_ref/avx512/test-backend-ops ]; then D=oracle/_ref/avx512; fi
for op in RMS_NORM ADD SUB MUL DIV SCALE SILU GELU GELU_QUICK GELU_ERF RELU SIGMOID TANH NEG ABS EXP REGLU GEGLU SWIGLU SWIGLU_OAI GEGLU_ERF GEGLU_QUICK SOFT_MAX ROPE CPY CONT DUP SET_ROWS ADD_ID ARGSORT; do
  GGML_BACKEND_PATH=$PWD/llama.cpp-gfx906_b200/lib/libggml-b200.so timeout 300 $D/test-backend-ops test -o $op > gpurun_out/tbo_$op.txt 2>&1
  rc=$?
  sed -i 's/\x1b\[[0-9;]*m//g' gpurun_out/tbo_$op.txt
  echo "$op rc=$rc OK $(grep -c 'OK$' gpurun_out/tbo_$op.txt) unsupported $(grep -c 'not supported' gpurun_out/tbo_$op.txt) FAIL $(grep -c 'FAIL' gpurun_out/tbo_$op.txt)"
done
grep -h FAIL gpurun_out/tbo_*.txt | head -40
