#!/bin/bash
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_mmq.py tests/test_gpu_parity.py -q -x -k "mxfp4" > gpurun_out/pytest_mxf.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_mxf.log
for nomxf in 0 1; do
  echo "== B200Q_NO_MXF=$nomxf"
  if [ $nomxf = 1 ]; then export B200Q_NO_MXF=1; fi
  timeout 120 python tools/moe_bench.py 2>&1 | tail -2
  for a in "mxfp4 2880 2880 512" "mxfp4 11520 2880 512" "mxfp4 2880 2880 64" "mxfp4 4096 4096 512" "mxfp4 2880 2880 32" "mxfp4 2880 2880 128" "mxfp4 11520 2880 128"; do timeout 60 python tools/mmq_one.py $a 20 2>&1 | tail -1; done
done
