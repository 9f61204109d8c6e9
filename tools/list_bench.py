#!/usr/bin/env python3
"""time the persistent decode list (b200q_list_*) on a workload's matmul list: whole pass, with / without dependency
barrier points, against the chain of grouped launches (gemv_kq_kernel) it replaces.
usage: tools/list_bench.py [8b|70b] [layers]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "llama.cpp-gfx906_b200"), os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)
import torch  # noqa: E402

import b200q  # noqa: E402
from b200q import workloads  # noqa: E402


def main():
    which = sys.argv[1] if len(sys.argv) > 1 else "8b"
    layers = int(sys.argv[2]) if len(sys.argv) > 2 else None
    ops = workloads.llama3_8b("q4_k_m", layers=layers) if which == "8b" else workloads.llama3_70b("q4_k", layers=layers)
    shards = [workloads.synth_weights_device(o, seed=1234 + i) for i, o in enumerate(ops)]
    acts = [torch.rand(o.k, device="cuda") * 2 - 1 for o in ops]
    outs = [torch.zeros(o.m, device="cuda") for o in ops]
    algo = sum(o.algo_bytes(1) for o in ops)
    SHARED = ("attn_k", "attn_v", "ffn_up")

    def build(dep):
        lo = []
        for i, o in enumerate(ops):
            same = o.name.split(".")[-1] in SHARED
            a = acts[i - 1] if same and i > 0 and ops[i - 1].k == o.k else acts[i]
            if same and i > 0:
                acts[i] = a
            lo.append((shards[i], a, outs[i], dep and not same and i > 0))
        return b200q.MatmulList(lo)

    def timeit(fn, reps=20, warm=3):
        for _ in range(warm):
            fn()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):
            fn()
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / reps

    print(f"{which}: {len(ops)} matmuls, {algo / 1e9:.3f} GB algorithmic per pass")
    for dep in (False, True):
        lst = build(dep)
        ms = timeit(lst.launch)
        print(f"list dep={int(dep)}: launches={lst.n_launches} barrier_points={lst.n_barrier_points} gens={lst.n_generations} "
              f"slots={lst.n_slots}  {ms:.4f} ms/pass  {algo / ms / 1e6:.1f} GB/s  ({algo / ms / 1e6 / 6540.8 * 100:.1f} % of measured HBM)")
        ref = [o.clone() for o in outs]
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            lst.launch()
        ms = timeit(g.replay)
        print(f"   graph replay: {ms:.4f} ms/pass  {algo / ms / 1e6:.1f} GB/s")
        del ref
    # the chain of grouped launches of round 1
    def chain():
        i = 0
        while i < len(ops):
            j = i + 1
            while j < len(ops) and j - i < 4 and ops[j].name.split(".")[-1] in SHARED and ops[j].qtype == ops[i].qtype and ops[j].k == ops[i].k:
                j += 1
            b200q.mul_mat_group(shards[i:j], acts[i].view(1, -1), outs=[outs[q].view(1, -1) for q in range(i, j)])
            i = j
    g = torch.cuda.CUDAGraph()
    chain()
    torch.cuda.synchronize()
    with torch.cuda.graph(g):
        chain()
    ms = timeit(g.replay)
    print(f"chain of grouped launches (graph): {ms:.4f} ms/pass  {algo / ms / 1e6:.1f} GB/s")


if __name__ == "__main__":
    main()
