#!/usr/bin/env python3
"""event trace of the persistent list kernel, CTA 0 (b200q_debug_list_stamps): where a warp's cycles go.
usage: tools/list_timeline.py [dep=1] [layers=4]"""
import ctypes as C
import os
import sys
from collections import defaultdict

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "llama.cpp-gfx906_b200"), os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)
import torch  # noqa: E402

import b200q  # noqa: E402
from b200q import workloads  # noqa: E402

dep = int(sys.argv[1]) if len(sys.argv) > 1 else 1
layers = int(sys.argv[2]) if len(sys.argv) > 2 else 4
NEV = 4096
ops = workloads.llama3_8b("q4_k_m", layers=layers, with_output=False)
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
for _ in range(3):
    lst.launch()
torch.cuda.synchronize()
ts = torch.zeros(8 * NEV * 2, dtype=torch.int64, device="cuda")
L = b200q.lib()
L.b200q_debug_list_stamps.argtypes = [C.c_void_p]
L.b200q_debug_list_stamps(C.c_void_p(ts.data_ptr()))
lst.launch()
torch.cuda.synchronize()
L.b200q_debug_list_stamps(None)
t = ts.cpu().view(8, NEV, 2).numpy()
NAMES = {1: "enter_switch", 2: "bp_done", 3: "quantized", 4: "switch_barrier", 5: "loop_top", 6: "stage_landed", 7: "computed",
         9: "refill_issued", 8: "stored"}
MHZ = 1965.0
print(f"dep={dep} layers={layers} gens={lst.n_generations}  (cycles at {MHZ:.0f} MHz)")
for w in (0, 5):
    ev = [(int(a) >> 32, int(a) & 0xffffffff, int(c)) for a, c in t[w] if a != 0 or c != 0]
    sums, cnts = defaultdict(int), defaultdict(int)
    for (pt, px, pc), (tg, x, c) in zip(ev[:-1], ev[1:]):
        sums[(pt, tg)] += c - pc
        cnts[(pt, tg)] += 1
    total = ev[-1][2] - ev[0][2]
    print(f"warp {w}: {len(ev)} events, {total} cycles = {total / MHZ:.1f} us")
    for k in sorted(sums, key=lambda k: -sums[k]):
        print(f"   {str(NAMES.get(k[0], k[0])):>15s} -> {str(NAMES.get(k[1], k[1])):<15s} n={cnts[k]:5d}  sum={sums[k] / MHZ:8.1f} us  avg={sums[k] / cnts[k]:8.0f} cyc")
    if w == 0:
        print("   first events of warp 0:")
        for (pt, px, pc), (tg, x, c) in list(zip(ev[:-1], ev[1:]))[:70]:
            print(f"      +{c - pc:7d}  {NAMES.get(tg, tg)} ({x})")
