import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "llama.cpp-gfx906_b200"), os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)
import torch, b200q, refs
from b200q import workloads
t = {"q4_K": refs.Q4_K, "q6_K": refs.Q6_K}[sys.argv[1] if len(sys.argv) > 1 else "q4_K"]
k = int(sys.argv[2]) if len(sys.argv) > 2 else 4096
op = workloads.MatMul("x", t, k, 4096)
w = workloads.synth_weights_device(op, seed=1)
a = torch.rand(k, device="cuda") * 2 - 1
outs = [torch.zeros(4096, device="cuda") for _ in range(32)]
lst = b200q.MatmulList([(w, a, o, False) for o in outs])
for _ in range(4):
    lst.launch()
torch.cuda.synchronize()
