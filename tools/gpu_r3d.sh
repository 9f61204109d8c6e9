#!/bin/bash
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_parity.py tests/test_gpu_plugin.py -x -q > gpurun_out/pytest_q5k.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/pytest_q5k.log
python - <<'PY'
import sys, os
sys.path.insert(0, 'llama.cpp-gfx906_b200'); sys.path.insert(0, 'tests')
import numpy as np, torch, b200q, refs
for t in (refs.Q5_K, refs.Q4_K):
    for (m, k) in ((4096, 4096), (14336, 4096), (4096, 14336)):
        base = refs.random_blocks(t, 512, k, seed=1)
        ws = [b200q.upload(t, (m, k), np.tile(base, ((m + 511) // 512, 1))[:m]) for _ in range(8)]
        act = torch.rand(1, k, device="cuda") * 2 - 1
        out = torch.empty(1, m, device="cuda")
        for w in ws: b200q.mul_mat(w, act, out=out)
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            for _ in range(5):
                for w in ws: b200q.mul_mat(w, act, out=out)
        g.replay(); torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); g.replay(); e1.record(); torch.cuda.synchronize()
        us = e0.elapsed_time(e1) * 1e3 / 40
        by = m * refs.row_size(t, k)
        print(f"{refs.NAMES[t]} {m}x{k} n=1: {us:.2f} us per launch, {by / us / 1e3:.0f} GB/s")
PY
