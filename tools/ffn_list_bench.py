"""ffn_gate / ffn_up / SWIGLU / ffn_down of 32 Llama-3-8B layers (distinct weights: 3.2 GB, nothing stays in L2) as
(a) list [gate, up] + a GLU kernel + list [down] per layer, (b) ONE list per layer with B200Q_LIST_SWIGLU, (c) one list for all
layers' gate/up/down (no GLU between: upper bound of the streaming) — CUDA-graph replays timed with events."""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))
sys.path.insert(0, os.path.join(ROOT, "llama.cpp-gfx906_b200"))
import b200q  # noqa: E402
import refs  # noqa: E402

L, dim, ff = 32, 4096, 14336
t = refs.Q4_K
rows = refs.make_weights(t, 512, dim, seed=1)
rows_d = refs.make_weights(t, 512, ff, seed=2)


def mk(m, k, base):
    w = np.tile(base, ((m + 511) // 512, 1))[:m]
    return b200q.upload(t, (m, k), w)


layers = [(mk(ff, dim, rows), mk(ff, dim, rows), mk(dim, ff, rows_d)) for _ in range(L)]
x = torch.from_numpy(refs.uniform_weights(1, dim, seed=3)).cuda().view(-1)
g = [torch.zeros(ff, device="cuda") for _ in range(L)]
u = [torch.zeros(ff, device="cuda") for _ in range(L)]
h = [torch.zeros(ff, device="cuda") for _ in range(L)]
o = [torch.zeros(dim, device="cuda") for _ in range(L)]
sep = [(b200q.MatmulList([(wg, x, g[i], False), (wu, x, u[i], False)]), b200q.MatmulList([(wd, h[i], o[i], False)])) for i, (wg, wu, wd) in enumerate(layers)]
fused = [b200q.MatmulList([(wg, x, g[i], False), (wu, x, u[i], False), (wd, g[i], o[i], True, u[i])]) for i, (wg, wu, wd) in enumerate(layers)]
plain = [b200q.MatmulList([(wg, x, g[i], False), (wu, x, u[i], False), (wd, g[i], o[i], True)]) for i, (wg, wu, wd) in enumerate(layers)]


def run_sep():
    for i in range(L):
        sep[i][0].launch()
        torch.mul(torch.nn.functional.silu(g[i]), u[i], out=h[i])
        sep[i][1].launch()


def run_fused():
    for i in range(L):
        fused[i].launch()


def run_plain():
    for i in range(L):
        plain[i].launch()


def timeit(fn, name):
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        fn()
        torch.cuda.synchronize()
        gr = torch.cuda.CUDAGraph()
        with torch.cuda.graph(gr, stream=s):
            fn()
        for _ in range(3):
            gr.replay()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(s)
        for _ in range(20):
            gr.replay()
        e1.record(s)
        torch.cuda.synchronize()
        us = e0.elapsed_time(e1) * 1000 / 20 / L
    by = (2 * ff * refs.row_size(t, dim) + dim * refs.row_size(t, ff)) / 1e3
    print(f"{name:28s} {us:7.2f} us per layer  ({by / us / 1e3:.2f} TB/s)")


timeit(run_sep, "list + glu(torch x2) + list")
timeit(run_fused, "one list, SWIGLU in quantizer")
timeit(run_plain, "one list, DEP only (no glu)")
