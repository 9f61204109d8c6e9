#!/bin/bash
# narrow-tile accumulator split of the tensor-core kernel: parity, then A/B timings (B200Q_MMQ_NO_ACCSPLIT=1 = before)
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_mmq.py tests/test_gpu_parity.py -x -q > gpurun_out/pytest_mmq.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_mmq.log
{
for a in "q4_K 14336 4096 32" "q4_K 14336 4096 64" "q8_0 4096 4096 32" "q4_K 4096 14336 32" "q4_K 4096 4096 17" "q6_K 4096 14336 48" "q4_0 4096 4096 64" "q8_0 1024 4096 32"; do
  echo -n "split  "; timeout 120 python tools/mmq_one.py $a 20 2>&1 | tail -1
  echo -n "before "; B200Q_MMQ_NO_ACCSPLIT=1 timeout 120 python tools/mmq_one.py $a 20 2>&1 | tail -1
done
echo "--- n = 512 (B ring in 32 KB slots; before: 46.6 / 63.9 / 106.5 / 39.0 / 42.6 us)"
for a in "q8_0 4096 4096 512" "q8_0 14336 4096 512" "q8_0 4096 14336 512" "q8_0 1024 4096 512" "q4_K 4096 4096 512" "q8_0 14336 4096 2048" "q8_0 128256 4096 512"; do
  timeout 120 python tools/mmq_one.py $a 20 2>&1 | tail -1
done
echo "--- MoE (mxfp4 goes through mmq_mxf; B200Q_NO_MXF=1 = the bf16 kernel)"
B200Q_NO_MXF=1 timeout 200 python tools/moe_bench.py 2>&1 | tail -3
B200Q_NO_MXF=1 B200Q_MMQ_NO_ACCSPLIT=1 timeout 200 python tools/moe_bench.py 2>&1 | tail -3
} > gpurun_out/accsplit.log 2>&1
cat gpurun_out/accsplit.log
