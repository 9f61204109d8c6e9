#!/bin/bash
# final GPU visit of the round: everything tools/gpu_round.sh collects + the reference's own test-backend-ops on the plugin
bash tools/gpu_round.sh
for op in MUL_MAT MUL_MAT_ID GET_ROWS; do
  timeout 600 bash tools/run_tbo.sh test -o $op > gpurun_out/tbo_$op.txt 2>&1; echo "tbo $op rc=$?"; tail -2 gpurun_out/tbo_$op.txt
done
tail -2 gpurun_out/pytest_gpu.log
