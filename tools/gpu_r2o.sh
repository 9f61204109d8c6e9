#!/bin/bash
# N-GPU bench only (torchrun) + row-split parity test
N=${1:-4}
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_rowsplit.py -q > gpurun_out/pytest_rowsplit_n$N.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_rowsplit_n$N.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 100 --warmup 3 > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err; echo "bench rc=$?"; tail -2 gpurun_out/bench_n$N.err; cut -c1-400 gpurun_out/bench_n$N.json
