"""quick GEMV timing sweep (development aid; bench.py is the contract).  Cold-HBM numbers: rotates over enough
distinct weight copies to exceed 4x L2."""
import sys, os, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "llama.cpp-gfx906_b200")); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np, torch
import b200q, refs

def bench(t, m, k, n, copies=None, iters=200, flags=b200q.FLAG_WEIGHTS_IMMUTABLE):
    rs = refs.row_size(t, k)
    wbytes = m * rs
    if copies is None:
        copies = max(2, int(600e6 // wbytes) + 1)
    base = refs.random_blocks(t, min(m, 512), k, seed=1)
    reps = (m + base.shape[0] - 1) // base.shape[0]
    host = np.tile(base, (reps, 1))[:m]
    ws = [b200q.upload(t, (m, k), host) for _ in range(copies)]
    act = torch.rand(n, k, device="cuda") * 2 - 1
    out = torch.empty(n, m, device="cuda")
    for i in range(3 * copies):
        b200q.mul_mat(ws[i % copies], act, out=out, flags=flags)
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for i in range(copies):
            b200q.mul_mat(ws[i], act, out=out, flags=flags)
    g.replay(); torch.cuda.synchronize()
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    reps = max(1, iters // copies)
    e0.record()
    for _ in range(reps):
        g.replay()
    e1.record(); torch.cuda.synchronize()
    us = e0.elapsed_time(e1) * 1e3 / (reps * copies)
    algo = wbytes + 4 * n * k + 4 * n * m
    print(f"{refs.NAMES[t]:6s} m={m:6d} k={k:6d} n={n} copies={copies:3d}: {us:8.2f} us  {algo / us / 1e3:8.1f} GB/s  ({algo/us/1e3/6547.8*100:5.1f}% of measured HBM)", flush=True)
    return us

if __name__ == "__main__":
    torch.cuda.set_device(0)
    which = sys.argv[1] if len(sys.argv) > 1 else "basic"
    if which == "basic":
        for (m, k) in [(4096, 4096), (14336, 4096), (4096, 14336), (1024, 4096), (128256, 4096)]:
            bench(refs.Q4_K, m, k, 1)
        bench(refs.Q4_0, 4096, 4096, 1)
        bench(refs.Q6_K, 128256, 4096, 1, copies=3)
        bench(refs.Q8_0, 14336, 4096, 1)
        bench(refs.MXFP4, 2880 * 4, 2880, 1)
        for n in (2, 4, 8):
            bench(refs.Q4_K, 14336, 4096, n)
        bench(refs.Q4_K, 14336, 4096, 1, flags=b200q.FLAG_NO_PDL)
    elif which == "cols":
        # 2..8 activation columns (gemv_mma.cu; B200Q_NO_MMA_GEMV=1 times the per-column kernel it replaces)
        types = [refs.Q4_K, refs.Q6_K] if len(sys.argv) < 3 else [{"q4_K": refs.Q4_K, "q5_K": refs.Q5_K, "q6_K": refs.Q6_K, "q4_0": refs.Q4_0, "q8_0": refs.Q8_0}[sys.argv[2]]]
        for t in types:
            for (m, k) in [(14336, 4096), (4096, 4096), (4096, 14336), (128256, 4096)]:
                for n in (2, 4, 8):
                    bench(t, m, k, n, copies=3 if m > 100000 else None)
    elif which == "mid":
        # 9..64 columns: B200Q_MMQ_MIN_N=9 -> tensor-core MMQ, B200Q_MMQ_MIN_N=100000 -> passes of the 8-column MMA GEMV
        t = {"q4_K": refs.Q4_K, "q6_K": refs.Q6_K, "q8_0": refs.Q8_0, "q4_0": refs.Q4_0}[sys.argv[2]]
        for (m, k) in [(14336, 4096), (4096, 4096), (4096, 14336)]:
            for n in (9, 12, 16, 24, 32, 48, 64):
                bench(t, m, k, n, iters=60)
    elif which == "one":
        t = {"q4_K": refs.Q4_K, "q6_K": refs.Q6_K, "q8_0": refs.Q8_0, "q4_0": refs.Q4_0, "mxfp4": refs.MXFP4}[sys.argv[2]]
        m, k, n = int(sys.argv[3]), int(sys.argv[4]), int(sys.argv[5])
        bench(t, m, k, n, copies=4, iters=16)
    elif which == "ts":
        # per-CTA timeline of the GEMV kernel inside a back-to-back chain (ns relative to the CTA's start stamp)
        t = {"q4_K": refs.Q4_K, "q6_K": refs.Q6_K, "q8_0": refs.Q8_0}[sys.argv[2]]
        m, k, n = int(sys.argv[3]), int(sys.argv[4]), int(sys.argv[5])
        host = np.tile(refs.random_blocks(t, 256, k, seed=1), ((m + 255) // 256, 1))[:m]
        ws = [b200q.upload(t, (m, k), host) for _ in range(8)]
        act = torch.rand(n, k, device="cuda"); out = torch.empty(n, m, device="cuda")
        L = b200q.lib()
        bufs = [torch.zeros(148 * 8, dtype=torch.int64, device="cuda") for _ in range(8)]
        import ctypes
        def chain():
            for i in range(8):
                L.b200q_debug_timestamps(ctypes.c_void_p(bufs[i].data_ptr()))
                b200q.mul_mat(ws[i], act, out=out, flags=int(os.environ.get('QB_FLAGS', '1')))
        chain(); torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            chain()
        for rep in range(3):
            g.replay()
        torch.cuda.synchronize()
        L.b200q_debug_timestamps(None)
        T = torch.stack(bufs).cpu().numpy().reshape(8, 148, 8)
        t0 = T[:, :, 0].min()
        for i in range(3, 8):
            a = T[i]
            names = ["start", "pdl_wait", "quant", "first_full", "end", "act_loaded", "q_done_t0"]
            order = [0, 1, 5, 6, 2, 3, 4]
            print(f"kernel {i}: " + "  ".join(f"{names[j]}: med {int(np.median(a[:, j]) - t0):6d} max {int(a[:, j].max() - t0):6d}" for j in order))
    elif which == "layer":
        # one Llama-3-8B layer + chain, grouped (qkv, o, gate+up, down) vs ungrouped launches, cold HBM (distinct copies)
        t = refs.Q4_K
        def mk(m, k):
            host = np.tile(refs.random_blocks(t, 256, k, seed=1), ((m + 255) // 256, 1))[:m]
            return b200q.upload(t, (m, k), host)
        L = 24
        layers = [dict(q=mk(4096, 4096), k=mk(1024, 4096), v=mk(1024, 4096), o=mk(4096, 4096), g=mk(14336, 4096),
                       u=mk(14336, 4096), d=mk(4096, 14336)) for _ in range(L)]
        a4 = torch.rand(1, 4096, device="cuda") * 2 - 1
        a14 = torch.rand(1, 14336, device="cuda") * 2 - 1
        outs = {n: torch.empty(1, layers[0][n].m, device="cuda") for n in "qkvogud"}
        byt = sum(w.nbytes for w in layers[0].values())
        def run(grouped):
            for l in layers:
                if grouped:
                    b200q.mul_mat_group([l["q"], l["k"], l["v"]], a4, [outs["q"], outs["k"], outs["v"]])
                    b200q.mul_mat(l["o"], a4, out=outs["o"])
                    b200q.mul_mat_group([l["g"], l["u"]], a4, [outs["g"], outs["u"]])
                    b200q.mul_mat(l["d"], a14, out=outs["d"])
                else:
                    for n in "qkvogu":
                        b200q.mul_mat(l[n], a4, out=outs[n])
                    b200q.mul_mat(l["d"], a14, out=outs["d"])
        for grouped in (False, True):
            run(grouped); torch.cuda.synchronize()
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g):
                run(grouped)
            g.replay(); torch.cuda.synchronize()
            e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(5):
                g.replay()
            e1.record(); torch.cuda.synchronize()
            us = e0.elapsed_time(e1) * 1e3 / 5 / L
            print(f"layer (q4_K, 7 matmuls, {byt/1e6:.0f} MB) grouped={grouped}: {us:7.2f} us/layer  {byt/us/1e3:7.1f} GB/s ({byt/us/1e3/6545*100:.1f}% of measured HBM)", flush=True)
    elif which == "mmq":
        # prompt processing (n = 512): TFLOP/s = 2*m*n*k / time (test-backend-ops' definition, :3114-3117)
        import json
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))) if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else {}
        bf16 = peaks.get("bf16_tflops", 1590.0)
        n = int(sys.argv[2]) if len(sys.argv) > 2 else 512
        for (t, m, k) in [(refs.Q8_0, 4096, 4096), (refs.Q8_0, 1024, 4096), (refs.Q8_0, 14336, 4096), (refs.Q8_0, 4096, 14336),
                          (refs.Q4_K, 14336, 4096), (refs.Q6_K, 4096, 14336), (refs.Q4_0, 4096, 4096), (refs.MXFP4, 2880, 2880),
                          (refs.Q8_0, 128256, 4096)]:
            base = refs.random_blocks(t, min(m, 512), k, seed=1)
            host = np.tile(base, ((m + base.shape[0] - 1) // base.shape[0], 1))[:m]
            w = b200q.upload(t, (m, k), host)
            act = torch.rand(n, k, device="cuda") * 2 - 1
            out = torch.empty(n, m, device="cuda")
            for _ in range(3):
                b200q.mul_mat(w, act, out=out)
            torch.cuda.synchronize()
            e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
            iters = 20
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g):
                for _ in range(iters):
                    b200q.mul_mat(w, act, out=out)
            g.replay(); torch.cuda.synchronize()
            e0.record()
            g.replay()
            e1.record(); torch.cuda.synchronize()
            us = e0.elapsed_time(e1) * 1e3 / iters
            tf = 2.0 * m * n * k / us / 1e6
            print(f"{refs.NAMES[t]:6s} m={m:6d} k={k:6d} n={n}: {us:9.2f} us  {tf:8.1f} TFLOP/s  ({tf / bf16 * 100:5.1f}% of measured bf16 {bf16:.0f})", flush=True)
    elif which == "small":
        for (m, k) in [(4096, 4096), (14336, 4096), (4096, 14336), (1024, 4096)]:
            bench(refs.Q4_K, m, k, 1)
