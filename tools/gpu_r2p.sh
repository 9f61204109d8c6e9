#!/bin/bash
# dense MUL_MAT (dense_mm.cu) + f16 activations: parity tests, then the reference's MUL_MAT / MUL_MAT_ID sweep on the plugin
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_dense.py -x -q > gpurun_out/pytest_dense.log 2>&1; echo "dense rc=$?"; tail -5 gpurun_out/pytest_dense.log
D=oracle/_ref; if grep -q avx512_vnni /proc/cpuinfo && [ -x oracle/_ref/avx512/test-backend-ops ]; then D=oracle/_ref/avx512; fi
for op in MUL_MAT MUL_MAT_ID GET_ROWS; do
  GGML_BACKEND_PATH=$PWD/llama.cpp-gfx906_b200/lib/libggml-b200.so timeout 900 $D/test-backend-ops test -o $op > gpurun_out/tbo_$op.txt 2>&1
  rc=$?
  sed -i 's/\x1b\[[0-9;]*m//g' gpurun_out/tbo_$op.txt
  echo "$op rc=$rc OK $(grep -c 'OK$' gpurun_out/tbo_$op.txt) unsupported $(grep -c 'not supported' gpurun_out/tbo_$op.txt) FAIL $(grep -c 'FAIL' gpurun_out/tbo_$op.txt)"
done
grep -h FAIL gpurun_out/tbo_MUL_MAT*.txt | head -40
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -3 gpurun_out/smoke.log
