#!/bin/bash
# compute-sanitizer over the multi-column decode kernel (memcheck: out-of-bounds / misaligned shared and global accesses;
# racecheck: shared-memory hazards around the rings, the team barrier and the cluster stores) on a few parity cases
mkdir -p gpurun_out
SEL='mma_columns and (shape3 or shape4 or shape6 or shape12) and (q4_K or q6_K or q8_0)'
timeout 600 compute-sanitizer --tool memcheck --error-exitcode 9 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "$SEL" > gpurun_out/sanitize_memcheck.txt 2>&1; echo "memcheck rc=$?"
tail -6 gpurun_out/sanitize_memcheck.txt
timeout 600 compute-sanitizer --tool racecheck --racecheck-report analysis python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "mma_columns and (shape3 or shape4) and (q4_K or q6_K)" > gpurun_out/sanitize_racecheck.txt 2>&1; echo "racecheck rc=$?"
tail -8 gpurun_out/sanitize_racecheck.txt
