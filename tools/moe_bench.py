"""gpt-oss-20B-shaped MoE MUL_MAT_ID (BASELINE configs[3]): 32 experts [2880 -> 2880] MXFP4, top-4, T = 1 and 512.
python tools/moe_bench.py"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "llama.cpp-gfx906_b200")); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np, torch
import b200q, refs

E, M, K, USED = 32, 2880, 2880, 4
base = refs.random_blocks(refs.MXFP4, 256, K, seed=5)
host = np.tile(base, ((E * M + 255) // 256, 1))[:E * M]
w = b200q.upload(refs.MXFP4, (E, M, K), host)
rs = b200q.row_size(refs.MXFP4, K)
for T in (1, 512):
    g = torch.Generator(device="cpu"); g.manual_seed(T)
    ids = torch.stack([torch.randperm(E, generator=g)[:USED] for _ in range(T)]).to(torch.int32).cuda()
    b = (torch.rand(T, 1, K, device="cuda") * 2 - 1)
    out = torch.empty(T, USED, M, device="cuda")
    b200q.mul_mat_id(w, b, ids, out=out); torch.cuda.synchronize()
    gr = torch.cuda.CUDAGraph()
    iters = 20 if T == 1 else 3
    with torch.cuda.graph(gr):
        for _ in range(iters):
            b200q.mul_mat_id(w, b, ids, out=out)
    gr.replay(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); gr.replay(); e1.record(); torch.cuda.synchronize()
    us = e0.elapsed_time(e1) * 1e3 / iters
    flops = 2.0 * M * K * T * USED
    wbytes = min(E, T * USED) * M * rs if T > 1 else USED * M * rs
    print(f"mxfp4 MoE 32x[{K}->{M}] top-{USED} T={T}: {us:.1f} us  {flops / us / 1e6:.2f} TFLOP/s  "
          f"distinct expert bytes {wbytes / 1e6:.1f} MB -> {wbytes / us / 1e3:.0f} GB/s")
