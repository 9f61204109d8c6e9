"""one MMQ problem, a few launches (ncu target): python tools/mmq_one.py q8_0 14336 4096 512 [iters]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "llama.cpp-gfx906_b200")); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np, torch
import b200q, refs
t = {"q4_K": refs.Q4_K, "q6_K": refs.Q6_K, "q8_0": refs.Q8_0, "q4_0": refs.Q4_0, "mxfp4": refs.MXFP4}[sys.argv[1]]
m, k, n = int(sys.argv[2]), int(sys.argv[3]), int(sys.argv[4])
iters = int(sys.argv[5]) if len(sys.argv) > 5 else 3
base = refs.random_blocks(t, min(m, 512), k, seed=1)
w = b200q.upload(t, (m, k), np.tile(base, ((m + 511) // 512, 1))[:m])
act = torch.rand(n, k, device="cuda") * 2 - 1
out = torch.empty(n, m, device="cuda")
for _ in range(iters):
    b200q.mul_mat(w, act, out=out)
torch.cuda.synchronize()
g = torch.cuda.CUDAGraph()
with torch.cuda.graph(g):
    for _ in range(iters):
        b200q.mul_mat(w, act, out=out)
g.replay(); torch.cuda.synchronize()
e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
e0.record()
g.replay()
e1.record(); torch.cuda.synchronize()
us = e0.elapsed_time(e1) * 1e3 / iters
print(f"{sys.argv[1]} m={m} k={k} n={n}: {us:.2f} us  {2.0 * m * n * k / us / 1e6:.1f} TFLOP/s")
