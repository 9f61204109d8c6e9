"""pp512 over the whole matmul set of a Llama-3-8B-shaped model (BASELINE configs[2]): every weight matrix of one layer
mix x n = 512 tokens through b200q.mul_mat (MMQ), CUDA-graph replay.  python tools/pp512_bench.py [q8_0|q4_k_m] [n]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "llama.cpp-gfx906_b200"))
import torch
import b200q
from b200q import workloads

mix = sys.argv[1] if len(sys.argv) > 1 else "q8_0"
n = int(sys.argv[2]) if len(sys.argv) > 2 else 512
ops = workloads.llama3_8b(mix)
# distinct weights per shape class are enough (the activations / results dominate L2 at n = 512 anyway): one tensor per
# (type, k, m), reused by every layer
cache, acts, outs = {}, {}, {}
for i, o in enumerate(ops):
    key = (o.qtype, o.k, o.m)
    if key not in cache:
        cache[key] = workloads.synth_weights_device(o, seed=77 + len(cache))
    if o.k not in acts:
        acts[o.k] = torch.rand(n, o.k, device="cuda") * 2 - 1
    if o.m not in outs:
        outs[o.m] = torch.empty(n, o.m, device="cuda")

def run():
    for o in ops:
        b200q.mul_mat(cache[(o.qtype, o.k, o.m)], acts[o.k], out=outs[o.m])

run(); torch.cuda.synchronize()
g = torch.cuda.CUDAGraph()
with torch.cuda.graph(g):
    run()
for _ in range(3):
    g.replay()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
reps = 5
e0.record()
for _ in range(reps):
    g.replay()
e1.record(); torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / reps
flops = sum(o.flops(n) for o in ops)
print(f"llama3-8b {mix} pp{n}: {len(ops)} matmuls, {ms:.3f} ms/batch, {flops / ms / 1e9:.1f} TFLOP/s, "
      f"{n / ms * 1e3:.0f} tok/s (matmuls only), {flops / ms / 1e9 / 1675 * 100:.1f}% of measured bf16 peak")
by = {}
