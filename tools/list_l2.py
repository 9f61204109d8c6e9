#!/usr/bin/env python3
"""consumer capability of the persistent list kernel: the SAME small matrix 32 times in one list (L2-resident weights, one
activation generation) — what the dot code + bookkeeping sustain when HBM is out of the way.  usage: tools/list_l2.py"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "llama.cpp-gfx906_b200"), os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)
import torch  # noqa: E402

import b200q  # noqa: E402
import refs  # noqa: E402
from b200q import workloads  # noqa: E402

for (t, m, k) in ((refs.Q4_K, 4096, 4096), (refs.Q6_K, 4096, 4096), (refs.Q4_K, 4096, 14336), (refs.Q6_K, 4096, 14336),
                  (refs.Q4_K, 8192, 8192), (refs.Q8_0, 4096, 4096), (refs.Q4_0, 4096, 4096), (refs.MXFP4, 2880, 2880)):
    op = workloads.MatMul("x", t, k, m)
    w = workloads.synth_weights_device(op, seed=1)
    a = torch.rand(k, device="cuda") * 2 - 1
    outs = [torch.zeros(m, device="cuda") for _ in range(32)]
    lst = b200q.MatmulList([(w, a, o, False) for o in outs])
    for _ in range(3):
        lst.launch()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10):
        lst.launch()
    e1.record()
    torch.cuda.synchronize()
    us = e0.elapsed_time(e1) * 1e3 / 10 / 32
    print(f"{refs.NAMES[t]:6s} {m}x{k}: {us:7.2f} us per matmul from L2 -> {op.weight_bytes / us / 1e3:8.1f} GB/s ({op.weight_bytes / us / 1e3 / 148:6.1f} GB/s per SM; HBM share is 44)")
