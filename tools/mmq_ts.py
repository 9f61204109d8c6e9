"""per-chunk pipeline timeline of CTA 0 of the MMQ kernel (clock64 stamps): python tools/mmq_ts.py q8_0 14336 4096 512"""
import os, sys, ctypes
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "llama.cpp-gfx906_b200")); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np, torch
import b200q, refs
t = {"q4_K": refs.Q4_K, "q6_K": refs.Q6_K, "q8_0": refs.Q8_0, "q4_0": refs.Q4_0, "mxfp4": refs.MXFP4}[sys.argv[1]]
m, k, n = int(sys.argv[2]), int(sys.argv[3]), int(sys.argv[4])
base = refs.random_blocks(t, min(m, 512), k, seed=1)
w = b200q.upload(t, (m, k), np.tile(base, ((m + 511) // 512, 1))[:m])
act = torch.rand(n, k, device="cuda") * 2 - 1
out = torch.empty(n, m, device="cuda")
for _ in range(3):
    b200q.mul_mat(w, act, out=out)
buf = torch.zeros(256 * 8, dtype=torch.int64, device="cuda")
L = b200q.lib()
L.b200q_debug_mmq_stamps(ctypes.c_void_p(buf.data_ptr()))
b200q.mul_mat(w, act, out=out)
torch.cuda.synchronize()
L.b200q_debug_mmq_stamps(None)
T = buf.cpu().numpy().reshape(256, 8)
t0 = T[0, 0]
ph = T[255]
print("phases (ns from kernel begin): setup done %d, mma last commit %d, acc ready seen by epilogue %d, epilogue done %d, all warps done %d" % tuple(int(ph[i] - ph[0]) for i in range(1, 6)))
names = ["mma:wait", "mma:A ok", "mma:B ok", "mma:commit", "dq:start", "dq:done", "dq:slot free", "B:free|mma:iss"]
print("chunk " + " ".join(f"{x:>12s}" for x in names))
for g in list(range(0, 20)) + list(range(56, 64)):
    print(f"{g:5d} " + " ".join(f"{int(T[g, j] - t0):12d}" for j in range(8)))
d = np.diff(T[8:60, 3])
print("cycles per chunk (MMA issue to issue), chunks 8..60: median", np.median(d), "mean", d.mean())
