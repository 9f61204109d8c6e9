#!/usr/bin/env python3
"""the Llama-3-8B Q4_K_M decode list (an n-layer slice, or the whole 225-matmul step with the output matrix), launched a few times
(ncu target): tools/list_one.py [dep=0] [layers=8] [with_output=0]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "llama.cpp-gfx906_b200"), os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)
import torch  # noqa: E402

import b200q  # noqa: E402
from b200q import workloads  # noqa: E402

dep = int(sys.argv[1]) if len(sys.argv) > 1 else 0
layers = int(sys.argv[2]) if len(sys.argv) > 2 else 8
with_output = bool(int(sys.argv[3])) if len(sys.argv) > 3 else False
ops = workloads.llama3_8b("q4_k_m", layers=layers, with_output=with_output)
shards = [workloads.synth_weights_device(o, seed=1234 + i) for i, o in enumerate(ops)]
acts = [torch.rand(o.k, device="cuda") * 2 - 1 for o in ops]
outs = [torch.zeros(o.m, device="cuda") for o in ops]
SHARED = ("attn_k", "attn_v", "ffn_up")
lo = []
for i, o in enumerate(ops):
    same = o.name.split(".")[-1] in SHARED and i > 0
    if same:
        acts[i] = acts[i - 1]
    lo.append((shards[i], acts[i], outs[i], bool(dep) and not same and i > 0))
lst = b200q.MatmulList(lo)
algo = sum(o.algo_bytes(1) for o in ops)
for _ in range(6):
    lst.launch()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
lst.launch()
e1.record()
torch.cuda.synchronize()
ms = e0.elapsed_time(e1)
print(f"{len(ops)} matmuls, {algo / 1e6:.1f} MB algorithmic, dep={dep}: {ms * 1e3:.1f} us, {algo / ms / 1e6:.1f} GB/s")
