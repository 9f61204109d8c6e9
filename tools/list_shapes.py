#!/usr/bin/env python3
"""steady-state rate of the persistent list kernel per (type, k, m): a list of ~600 MB of distinct copies of ONE shape,
no dependencies — shows which shapes are HBM-bound and which are issue/latency-bound inside the kernel.
usage: tools/list_shapes.py"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "llama.cpp-gfx906_b200"), os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)
import torch  # noqa: E402

import b200q  # noqa: E402
from b200q import workloads  # noqa: E402
from b200q.workloads import MatMul  # noqa: E402

SHAPES = [(b200q.Q4_K, 4096, 4096), (b200q.Q4_K, 4096, 1024), (b200q.Q4_K, 4096, 14336), (b200q.Q4_K, 14336, 4096),
          (b200q.Q6_K, 4096, 1024), (b200q.Q6_K, 14336, 4096), (b200q.Q6_K, 4096, 128256), (b200q.Q4_K, 4096, 128256),
          (b200q.Q4_0, 4096, 4096), (b200q.Q8_0, 4096, 4096), (b200q.MXFP4, 2880, 2880), (b200q.Q4_K, 8192, 8192),
          (b200q.Q4_K, 28672, 8192)]
if len(sys.argv) > 1:
    SHAPES = SHAPES[:int(sys.argv[1])]


def main():
    for t, k, m in SHAPES:
        op = MatMul("x", t, k, m)
        n = max(2, min(64, int(600e6 // op.weight_bytes)))
        ws = [workloads.synth_weights_device(op, seed=77 + i) for i in range(n)]
        act = torch.rand(k, device="cuda") * 2 - 1
        outs = [torch.zeros(m, device="cuda") for _ in range(n)]
        lst = b200q.MatmulList([(ws[i], act, outs[i], False) for i in range(n)])
        for _ in range(3):
            lst.launch()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        reps = 10
        e0.record()
        for _ in range(reps):
            lst.launch()
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / reps
        algo = n * op.algo_bytes(1)
        print(f"{b200q.TYPE_NAMES[t]:6s} {k:6d}->{m:6d} x{n:3d}  {ms * 1e3 / n:8.2f} us/matmul  {algo / ms / 1e6:7.1f} GB/s "
              f"({algo / ms / 1e6 / 6546.6 * 100:5.1f} %)  launches={lst.n_launches} slots={lst.n_slots}", flush=True)
        lst.close()
        del ws, outs, lst
        torch.cuda.empty_cache()


if __name__ == "__main__":
    main()
