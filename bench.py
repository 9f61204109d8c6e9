#!/usr/bin/env python3
"""bench.py — the measurement contract of this repo.

A "step" is ONE pass of the hot path over the whole matmul list of a decode token of the workload
(BASELINE.json configs[1]: Llama-3-8B shapes, Q4_K_M mix, n = 1 — 225 GEMVs, 4.6 GB of distinct quantized weights,
i.e. far larger than L2, so every step streams cold from HBM; no flush needed).

  value   algorithmic GB/s of the pass with everything resident in HBM (CUDA-graph replay of the 225 launches,
          CUDA events on the launching stream, max over ranks)
  e2e     the same pass through the C-ABI with HOST buffers: per step the activations are copied host->device from
          pinned memory, the pass runs, all results are copied device->host (copies inside the timed region; on one
          GPU the copies of a chunk of launches overlap the compute of its neighbours, B200Q_BENCH_E2E=serial = old)
  --impl reference   the reference's own ggml-cpu implementation (oracle/_ref, unmodified) of the same matmuls on
          the box's host cores, on a bounded sample of the workload

N > 1 (torchrun, one rank per GPU): the reference's row-split mode — every weight matrix is sharded by output rows,
each rank computes its slice, NCCL all-gathers the result slices (scaling = "strong": total work is fixed).
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
for p in (ROOT, os.path.join(ROOT, "llama.cpp-gfx906_b200"), os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)

METRIC = "llama3-8b Q4_K_M decode GEMV set (n=1): achieved HBM GB/s = algorithmic bytes / time"
UNIT = "GB/s"


def measured_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as f:
            d = json.load(f)
        return float(d.get("hbm_gbs", 6650.0)), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md)"


def measured_bf16():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as f:
            return float(json.load(f).get("bf16_tflops", 1675.0))
    return 1675.0


class ClockSampler:
    """samples nvidia-smi clocks / throttle reasons while the timed region runs"""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, device_index=0):
        self.rows = []
        self.proc = None
        self.dev = device_index

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-i", str(self.dev), "-lms", "100"], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        sm, mx, reasons = [], [], set()
        for r in self.rows:
            try:
                sm.append(float(r[1]))
                mx.append(float(r[2]))
            except (ValueError, IndexError):
                continue
            for name, col in (("hw_slowdown", 5), ("hw_thermal_slowdown", 6), ("sw_thermal_slowdown", 7),
                              ("sw_power_cap", 8)):
                if len(r) > col and r[col].lower().startswith("active"):
                    reasons.add(name)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="llama3-8b-q4_k_m", choices=["llama3-8b-q4_k_m", "llama3-70b-q4_k"])
    ap.add_argument("--cpu-seconds", type=float, default=12.0, help="budget of the cpu_baseline leg")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-llama-bench", action="store_true", help="skip the aux leg that runs the reference's llama-bench on the plugin")
    return ap.parse_args()


def workload_ops(name):
    from b200q import workloads
    if name == "llama3-8b-q4_k_m":
        return workloads.llama3_8b("q4_k_m")
    return workloads.llama3_70b("q4_k")


# --------------------------------------------------------------------------------------------------------------
# the reference's CPU implementation (oracle/_ref, unmodified ggml-cpu) — checker / baseline only
# --------------------------------------------------------------------------------------------------------------
def cpu_sample_ops(ops):
    """bounded sample of the workload: one 'more-bits' layer and one plain layer of the per-layer list (14 matmuls,
    every distinct shape/type of the layers; the 431 MB output head is left out to bound the run time)"""
    names = {}
    for o in ops:
        if o.name.startswith("blk.0.") or o.name.startswith("blk.4."):
            names[o.name] = o
    return list(names.values())


def run_cpu_reference(ops, seconds, threads=0):
    """returns dict(value GB/s, cores, kind, sample, ms_per_pass) for repeated passes over `ops` on the host cores"""
    import numpy as np
    import refs
    R = refs.ref()
    kind = "reference" if R is not None else "port"
    cores = os.cpu_count() or 1
    data = []
    for i, o in enumerate(ops):
        base = refs.random_blocks(o.qtype, min(o.m, 256), o.k, seed=100 + i)
        w = np.ascontiguousarray(np.tile(base, ((o.m + base.shape[0] - 1) // base.shape[0], 1))[:o.m])
        act = refs.uniform_weights(1, o.k, seed=200 + i)
        data.append((o, w, act))
    algo = sum(o.algo_bytes(1) for o in ops)

    def one_pass():
        t = 0.0
        for o, w, act in data:
            if R is not None:
                _, secs = refs.ref_mul_mat(o.qtype, w, act, threads=threads, repeat=1)
            else:
                t0 = time.perf_counter()
                refs.orc_mul_mat(o.qtype, w, act)
                secs = time.perf_counter() - t0
            t += secs
        return t

    one_pass()  # warm-up (thread pool, page faults)
    times = []
    t_start = time.perf_counter()
    while time.perf_counter() - t_start < seconds or len(times) < 2:
        times.append(one_pass())
        if len(times) >= 200:
            break
    best = sorted(times)[len(times) // 2]
    feats = R.ref_cpu_features().decode() if R is not None else "scalar C port (oracle/oracle.c)"
    return {"value": algo / best / 1e9, "unit": UNIT, "cores": cores if kind == "reference" else 1, "kind": kind,
            "sample": f"{len(ops)} matmuls of layers 0 and 4 ({algo / 1e6:.0f} MB algorithmic bytes/pass), "
                      f"median of {len(times)} passes; {feats}",
            "ms_per_pass": best * 1e3, "passes": len(times)}


def main_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    ops = workload_ops(args.workload)
    sample = cpu_sample_ops(ops)
    # each "step" is one pass over the bounded sample; W warm-up passes, K timed passes
    import numpy as np
    import refs
    R = refs.ref()
    data = []
    for i, o in enumerate(sample):
        base = refs.random_blocks(o.qtype, min(o.m, 256), o.k, seed=100 + i)
        w = np.ascontiguousarray(np.tile(base, ((o.m + base.shape[0] - 1) // base.shape[0], 1))[:o.m])
        data.append((o, w, refs.uniform_weights(1, o.k, seed=200 + i)))
    algo = sum(o.algo_bytes(1) for o in sample)

    def step():
        t = 0.0
        for o, w, act in data:
            if R is not None:
                t += refs.ref_mul_mat(o.qtype, w, act, threads=0, repeat=1)[1]
            else:
                t0 = time.perf_counter()
                refs.orc_mul_mat(o.qtype, w, act)
                t += time.perf_counter() - t0
        return t

    for _ in range(max(1, args.warmup)):
        step()
    total = sum(step() for _ in range(args.steps))
    value = algo * args.steps / total / 1e9
    cores = (os.cpu_count() or 1) if R is not None else 1
    kind = "reference" if R is not None else "port"
    sample_txt = f"{len(sample)} matmuls of layers 0 and 4 of the workload ({algo / 1e6:.0f} MB/step), all host threads"
    out = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": total / args.steps * 1e3, "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "int8 dot / f32 accumulate (ggml-cpu vec_dot)", "data": "synthetic",
        "config": {"workload": args.workload + " decode GEMV set, n=1 (bounded sample)", "sample": sample_txt},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": kind, "sample": sample_txt},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(out))
    return 0


# --------------------------------------------------------------------------------------------------------------
# the B200 arm
# --------------------------------------------------------------------------------------------------------------
SHARED_LEAVES = ("attn_k", "attn_v", "ffn_up")  # read the same vector as the op before them (wq/wk/wv, ffn_gate/ffn_up)


def _timeit(torch, fn, reps, warm=3):
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


def aux_single_gpu(torch, b200q, workloads, peak_hbm, bf16_peak):
    """BASELINE configs 1, 3, 4 — not the headline; each entry: us, achieved, unit, frac, peak"""
    import refs
    aux = {}
    # config 1: Q4_0 4096 x 4096, n = 1; 64 distinct weight copies (604 MB > 4 x L2) so every launch streams from HBM
    try:
        op = workloads.MatMul("c1", refs.Q4_0, 4096, 4096)
        ws = [workloads.synth_weights_device(op, seed=900 + i) for i in range(64)]
        a = torch.rand(4096, device="cuda") * 2 - 1
        outs = [torch.zeros(4096, device="cuda") for _ in ws]
        lst = b200q.MatmulList([(w, a, o, False) for w, o in zip(ws, outs)])
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            lst.launch()
        us_list = _timeit(torch, g.replay, 20) * 1e3 / len(ws)
        g2 = torch.cuda.CUDAGraph()
        b200q.mul_mat(ws[0], a.view(1, -1), out=outs[0].view(1, -1))
        torch.cuda.synchronize()
        with torch.cuda.graph(g2):
            for w, o in zip(ws, outs):
                b200q.mul_mat(w, a.view(1, -1), out=o.view(1, -1))
        us_chain = _timeit(torch, g2.replay, 20) * 1e3 / len(ws)
        bytes_ = op.algo_bytes(1)
        aux["config1_q4_0_4096x4096_n1"] = {
            "us": us_list, "achieved": bytes_ / us_list / 1e3, "unit": "GB/s", "frac": bytes_ / us_list / 1e3 / peak_hbm, "peak": peak_hbm,
            "how": "64 distinct weight copies (604 MB, > 4 x L2) as one persistent list launch, us per matmul",
            "per_launch_chain_us": us_chain, "per_launch_chain_frac": bytes_ / us_chain / 1e3 / peak_hbm}
        del ws, outs, lst, g, g2
    except Exception as e:
        aux["config1_q4_0_4096x4096_n1"] = {"error": repr(e)}
    # config 3: Q8_0 pp512 over the matmul set of 4 layers of Llama-3-8B shapes (928 MB of weights > L2)
    try:
        n = 512
        ops = workloads.llama3_8b("q8_0", layers=4, with_output=False)
        ws = [workloads.synth_weights_device(o, seed=950 + i) for i, o in enumerate(ops)]
        acts = {k: torch.rand(n, k, device="cuda") * 2 - 1 for k in {o.k for o in ops}}
        outs = {m: torch.empty(n, m, device="cuda") for m in {o.m for o in ops}}

        def run():
            for o, w in zip(ops, ws):
                b200q.mul_mat(w, acts[o.k], out=outs[o.m])
        run()
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            run()
        ms = _timeit(torch, g.replay, 10)
        flops = sum(o.flops(n) for o in ops)
        tf = flops / ms / 1e9
        per = {}
        for o, w in zip(ops[:7], ws[:7]):
            us = _timeit(torch, lambda: b200q.mul_mat(w, acts[o.k], out=outs[o.m]), 10) * 1e3
            per[f"{o.k}->{o.m}"] = {"us": us, "tflops": o.flops(n) / us / 1e6, "frac_bf16": o.flops(n) / us / 1e6 / bf16_peak}
        aux["config3_q8_0_pp512"] = {
            "us": ms * 1e3, "achieved": tf, "unit": "TFLOP/s", "frac": tf / bf16_peak, "peak": bf16_peak,
            "frac_of_nominal_int8_4500": tf / 4500.0, "tok_s_matmul_only_32_layers": n / (ms * 8) * 1e3,
            "how": "28 matmuls (4 layers, 928 MB of Q8_0 weights) x 512 tokens, CUDA-graph replay; tcgen05 kind::f16 on bf16-dequantized tiles",
            "per_shape_L2_warm": per}
        del ws, acts, outs, g
    except Exception as e:
        aux["config3_q8_0_pp512"] = {"error": repr(e)}
    # config 4: gpt-oss-20B-shaped MoE, 32 experts [2880 -> 2880] MXFP4, top-4, T = 1 and T = 512
    try:
        import numpy as np
        E, M, K, USED = 32, 2880, 2880, 4
        base = refs.random_blocks(refs.MXFP4, 256, K, seed=5)
        host = np.tile(base, ((E * M + 255) // 256, 1))[:E * M]
        nrep = 5  # 5 expert tensors of 141 MB rotate (705 MB > 4 x L2)
        wsx = [b200q.upload(refs.MXFP4, (E, M, K), host) for _ in range(nrep)]
        rs = b200q.row_size(refs.MXFP4, K)
        for T in (1, 512):
            gen = torch.Generator(device="cpu")
            gen.manual_seed(T)
            ids = torch.stack([torch.randperm(E, generator=gen)[:USED] for _ in range(T)]).to(torch.int32).cuda()
            b = torch.rand(T, 1, K, device="cuda") * 2 - 1
            out = torch.empty(T, USED, M, device="cuda")
            b200q.mul_mat_id(wsx[0], b, ids, out=out)
            torch.cuda.synchronize()
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g):
                for w in wsx:
                    b200q.mul_mat_id(w, b, ids, out=out)
            us = _timeit(torch, g.replay, 10 if T == 1 else 4) * 1e3 / nrep
            if T == 1:
                by = USED * M * rs + K * 4 + USED * M * 4
                aux["config4_mxfp4_moe_T1"] = {"us": us, "achieved": by / us / 1e3, "unit": "GB/s", "frac": by / us / 1e3 / peak_hbm,
                                               "peak": peak_hbm, "how": "4 of 32 experts [2880->2880], 5 rotating expert tensors"}
            else:
                fl = 2.0 * M * K * T * USED
                aux["config4_mxfp4_moe_T512"] = {"us": us, "achieved": fl / us / 1e6, "unit": "TFLOP/s", "frac": fl / us / 1e6 / bf16_peak,
                                                 "peak": bf16_peak, "how": "2048 (token, slot) rows over 32 experts, native block-scaled tcgen05 kind::mxf8f6f4 (hi + lo e4m3 activations); frac is against the measured bf16 peak"}
        del wsx
    except Exception as e:
        aux["config4_mxfp4_moe"] = {"error": repr(e)}
    return aux


def aux_llama_bench():
    """tg128 / pp512 tok/s of the reference's OWN llama-bench (oracle/_ref/llama-bench, compiled unmodified by `make -C oracle
    llama`; the second half of BASELINE.json's metric) on the plugin, synthetic-weight Llama-3-8B Q4_K_M GGUF written by
    tools/make_synth_gguf.py.  Runs in a child process after this process has released nothing it needs (the model is 4.9 GB);
    absent binary -> a note, never an error for the headline."""
    import re
    import subprocess
    import tempfile
    d = os.path.join(ROOT, "oracle", "_ref")
    try:
        with open("/proc/cpuinfo") as f:
            if "avx512_vnni" in f.read() and os.path.exists(os.path.join(d, "avx512", "llama-bench")):
                d = os.path.join(d, "avx512")
    except OSError:
        pass
    exe = os.path.join(d, "llama-bench")
    plugin = os.path.join(ROOT, "llama.cpp-gfx906_b200", "lib", "libggml-b200.so")
    if not (os.path.exists(exe) and os.path.exists(plugin)):
        return {"unavailable": "oracle/_ref/llama-bench not built (make -C oracle llama)"}
    gguf = os.path.join(tempfile.gettempdir(), f"b200q-synth-{os.getpid()}.gguf")
    try:
        subprocess.run([sys.executable, os.path.join(ROOT, "tools", "make_synth_gguf.py"), "--out", gguf, "--layers", "32", "--ftype", "q4_k_m"],
                       check=True, capture_output=True, timeout=300)
        env = dict(os.environ, GGML_BACKEND_PATH=plugin)
        r = subprocess.run([exe, "-m", gguf, "-ngl", "99", "-fa", "0", "-p", "512", "-n", "128", "-r", "3"], env=env, capture_output=True,
                           text=True, timeout=600)
        res = {}
        for line in r.stdout.splitlines():
            m = re.search(r"\|\s*(B200|CPU)\s*\|.*\|\s*(pp512|tg128)\s*\|\s*([0-9.]+)\s*±\s*([0-9.]+)", line)
            if m:
                res[m.group(2)] = {"tok_s": float(m.group(3)), "stdev": float(m.group(4)), "backend": m.group(1)}
        if not res:
            return {"error": (r.stdout + r.stderr)[-400:]}
        res["how"] = ("reference llama-bench (unmodified) -ngl 99 -fa 0 -p 512 -n 128 -r 3 through GGML_BACKEND_PATH=libggml-b200.so; "
                      "synthetic-weight Llama-3-8B Q4_K_M GGUF (4.9 GB, no tokenizer)")
        return res
    except Exception as e:  # noqa: BLE001
        return {"error": repr(e)}
    finally:
        try:
            os.remove(gguf)
        except OSError:
            pass


def main_b200(args):
    import ctypes as C

    import torch

    import b200q
    from b200q import workloads

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    # stdout carries exactly ONE JSON line: whatever libraries print meanwhile ("NCCL version ...") goes to stderr
    sys.stdout.flush()
    saved_stdout = os.dup(1)
    os.dup2(2, 1)
    assert torch.cuda.is_available(), "bench.py needs a GPU (no CPU fallback)"
    torch.cuda.set_device(local)
    dist = None
    if world > 1:
        import torch.distributed as dist
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    L = b200q.lib()
    from b200q import rowsplit

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    def reduce_max(x):
        if dist is None:
            return x
        t = torch.tensor([x], device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    class Pass:
        """the resident state of one workload: this rank's row shard of every matrix, all activation vectors, results"""

        def __init__(self, name, seed0=1234):
            self.ops = workload_ops(name)
            ops = self.ops
            self.shards = []
            for i, o in enumerate(ops):
                parts = rowsplit.row_split(o.m, world)
                row0, nrows = parts[rank]
                shard = workloads.synth_weights_device(o, seed=seed0 + i, rows=nrows, row0=row0)
                self.shards.append(rowsplit.RowSplitWeight(o.qtype, o.m, o.k, rank, world, parts, shard))
            self.k_off, self.m_off = [0], [0]
            for o in ops:
                self.k_off.append(self.k_off[-1] + o.k)
                self.m_off.append(self.m_off[-1] + o.m)
            g = torch.Generator(device="cuda")
            g.manual_seed(4321)
            self.act_dev = torch.rand(self.k_off[-1], device="cuda", generator=g) * 2 - 1
            self.algo = sum(o.algo_bytes(1) for o in ops)
            # results: two buffers that alternate by step parity (row split: a peer may still read step n while step n + 1
            # is stored into the other one; include/b200q.h, b200q_list_peers)
            self.peer = None
            if world > 1:
                self.peer = [rowsplit.PeerDst(self.m_off[-1], rank, world) for _ in range(2)]
                self.out = [p.tensor for p in self.peer]
                self.sync = rowsplit.PeerSync(rank, world)
            else:
                self.out = [torch.zeros(self.m_off[-1], device="cuda") for _ in range(2)]
            for o in self.out:
                o.zero_()

        def same_vec(self, i):
            o, ops = self.ops[i], self.ops
            return i > 0 and o.name.split(".")[-1] in SHARED_LEAVES and ops[i - 1].k == o.k

        def list_ops(self, buf, lo, hi, dep):
            """b200q_list ops of matmuls [lo, hi): consecutive ops of one group read the same vector; with dep every NEW vector
            waits for all earlier ops (the model's chain: qkv -> wo -> gate/up -> down -> next layer)"""
            res, src = [], None
            for i in range(lo, hi):
                o = self.ops[i]
                if not (self.same_vec(i) and src is not None):
                    src = self.act_dev[self.k_off[i]:self.k_off[i + 1]]
                sh = self.shards[i]
                dptr = (self.peer[buf].addr(self.m_off[i] + sh.row0) if self.peer else
                        self.out[buf][self.m_off[i]:self.m_off[i + 1]])
                w = sh.shard if sh.nrows > 0 else b200q.QTensor(o.qtype, (0, o.k), self.act_dev)
                res.append((w, src, dptr, bool(dep and i > lo and not self.same_vec(i))))
            return res

        def make_list(self, buf, lo=0, hi=None, dep=True):
            hi = len(self.ops) if hi is None else hi
            peers = None
            if self.peer:
                peers = (rank, world, self.peer[buf].deltas, self.sync.ptrs)
            return b200q.MatmulList(self.list_ops(buf, lo, hi, dep), peers=peers)

    P = Pass(args.workload)
    ops = P.ops
    # One GPU: the synthetic matmuls read their own pre-resident vectors, nothing in the workload depends on anything — the
    # list runs without barrier points (the dependent chain is reported in aux).  Row split: a matmul group's rows come from
    # all GPUs, so the headline honours the model's chain (every new vector waits for all earlier groups on all GPUs).
    dep_chain = os.environ.get("B200Q_BENCH_DEP", "1" if world > 1 else "0") != "0"
    lists = [P.make_list(b, dep=dep_chain) for b in range(2)]
    for l in lists:
        l.launch()
    torch.cuda.synchronize()
    barrier()
    graphs = []
    for l in lists:
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            l.launch()
        graphs.append(g)
    torch.cuda.synchronize()
    barrier()

    # ---- value: device-resident ---------------------------------------------------------------------------
    step_no = [0]

    def dev_step():
        graphs[step_no[0] & 1].replay()
        step_no[0] += 1

    for _ in range(max(4, args.warmup + (args.warmup & 1))):
        dev_step()
    barrier()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.steps):
        dev_step()
    e1.record()
    barrier()
    ms_dev = reduce_max(e0.elapsed_time(e1)) / args.steps
    clocks = sampler.stop() if rank == 0 else None
    if step_no[0] & 1:
        dev_step()  # keep the step parity of all ranks even
    barrier()

    # ---- row-split parity: gathered peer-store results == the un-sharded single-GPU result, bit for bit --------
    rowsplit_parity = None
    if world > 1:
        mism = 0
        seen = set()
        if rank == 0:
            for i, o in enumerate(ops):
                key = (o.qtype, o.k, o.m)
                if key in seen:
                    continue
                seen.add(key)
                # the same bytes every rank generated for its shard, assembled into the whole matrix on this GPU
                parts = P.shards[i].parts
                pieces = [workloads.synth_weights_device(o, seed=1234 + i, rows=nr, row0=r0) for (r0, nr) in parts if nr > 0]
                rs = b200q.row_size(o.qtype, o.k)
                full = torch.zeros(o.m * rs + 64, dtype=torch.uint8, device="cuda")
                pos = 0
                for pc in pieces:
                    nb = pc.m * rs
                    full[pos:pos + nb] = pc.data[:nb]
                    pos += nb
                dst = torch.zeros(o.m, device="cuda")
                src = P.act_dev[P.k_off[i]:P.k_off[i + 1]]
                if P.same_vec(i):
                    src = P.act_dev[P.k_off[i - 1]:P.k_off[i]] if not P.same_vec(i - 1) else P.act_dev[P.k_off[i - 2]:P.k_off[i - 1]]
                one = b200q.MatmulList([(b200q.QTensor(o.qtype, (o.m, o.k), full), src, dst, False)])
                one.launch()
                torch.cuda.synchronize()
                for b in range(2):
                    got = P.out[b][P.m_off[i]:P.m_off[i + 1]]
                    mism += int((got.view(torch.int32) != dst.view(torch.int32)).sum().item())
                del full, pieces, one
        rowsplit_parity = "bit-exact" if mism == 0 else f"{mism} mismatches"
        barrier()

    # ---- e2e: host buffers through the C-ABI, copies inside the timed region --------------------------------
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    act_host = P.act_dev.cpu().pin_memory()
    n_out = P.m_off[-1]
    # every rank brings in all activations (they are replicated); the results leave from rank 0's buffer only when the
    # job is a row split (every rank holds all rows after the kernel; one copy of them is the job's result)
    out_host = torch.empty(n_out, dtype=torch.float32).pin_memory()
    h2d = act_host.numel() * 4
    d2h = n_out * 4
    n_chunks = int(os.environ.get("B200Q_BENCH_E2E_CHUNKS", "8"))
    layer_starts = [i for i, o in enumerate(ops) if o.name.endswith("attn_q")] + [len(ops)]
    per = max(1, (len(layer_starts) - 1 + n_chunks - 1) // n_chunks)
    cuts = layer_starts[:-1:per] + [len(ops)]
    if cuts[-2] != len(ops) - 1 and ops[-1].name == "output":
        cuts = cuts[:-1] + [len(ops) - 1, len(ops)]  # the head is its own chunk
    chunk_lists = [[P.make_list(b, lo, hi, dep=dep_chain) for lo, hi in zip(cuts[:-1], cuts[1:])] for b in range(2)]
    s_in, s_out = torch.cuda.Stream(), torch.cuda.Stream()

    def copy(dst_t, src_t, lo, hi, kind, st):
        rc = L.b200q_memcpy(C.c_void_p(dst_t.data_ptr() + 4 * lo), C.c_void_p(src_t.data_ptr() + 4 * lo),
                            C.c_size_t(4 * (hi - lo)), kind, C.c_void_p(st.cuda_stream))
        assert rc == 0

    e2e_graphs = []
    for b in range(2):
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            main = torch.cuda.current_stream()
            s_in.wait_stream(main)
            s_out.wait_stream(main)
            # activations: the first chunk's slice, then all the rest (two joins); results leave chunk by chunk
            copy(P.act_dev, act_host, P.k_off[cuts[0]], P.k_off[cuts[1]], 1, s_in)
            ev0 = torch.cuda.Event()
            ev0.record(s_in)
            ev1 = None
            if len(cuts) > 2:
                copy(P.act_dev, act_host, P.k_off[cuts[1]], P.k_off[-1], 1, s_in)
                ev1 = torch.cuda.Event()
                ev1.record(s_in)
            for c, lst in enumerate(chunk_lists[b]):
                if c == 0:
                    main.wait_event(ev0)
                elif c == 1 and ev1 is not None:
                    main.wait_event(ev1)
                lst.launch()
                if rank == 0:
                    done = torch.cuda.Event()
                    done.record(main)
                    s_out.wait_event(done)
                    copy(out_host, P.out[b], P.m_off[cuts[c]], P.m_off[cuts[c + 1]], 2, s_out)
            main.wait_stream(s_in)
            main.wait_stream(s_out)
        e2e_graphs.append(g)
    torch.cuda.synchronize()
    barrier()
    e2e_no = [0]

    def e2e_step():
        e2e_graphs[e2e_no[0] & 1].replay()
        e2e_no[0] += 1
        assert L.b200q_stream_synchronize(stream) == 0

    # the pipelined graph must deliver what the resident pass computed
    out_host.zero_()
    e2e_step()
    e2e_step()
    torch.cuda.synchronize()
    e2e_ok = True
    if rank == 0:
        e2e_ok = bool(torch.equal(out_host, P.out[1].cpu())) and bool(torch.equal(P.out[0], P.out[1]))
    for _ in range(max(4, args.warmup + (args.warmup & 1))):
        e2e_step()
    barrier()
    t0 = time.perf_counter()
    e0.record()
    for _ in range(args.steps):
        e2e_step()
    e1.record()
    torch.cuda.synchronize()
    wall_ms = (time.perf_counter() - t0) * 1e3
    barrier()
    ms_e2e = reduce_max(max(e0.elapsed_time(e1), wall_ms)) / args.steps
    finite = bool(torch.isfinite(out_host).all())
    if e2e_no[0] & 1:
        e2e_step()
    barrier()

    # ---- row split, the other dependency policy (aux): the same step without / with the per-group barrier points ----------
    variant = None
    if world > 1:
        try:
            vl = [P.make_list(b, dep=not dep_chain) for b in range(2)]
            vn = [0]

            def vstep():
                vl[vn[0] & 1].launch()
                vn[0] += 1
            for _ in range(4):
                vstep()
            barrier()
            kv = max(4, min(args.steps, 40)) & ~1
            e0.record()
            for _ in range(kv):
                vstep()
            e1.record()
            barrier()
            msv = reduce_max(e0.elapsed_time(e1)) / kv
            variant = {("no_dependencies_upper_bound" if dep_chain else "with_dependencies"):
                       {"ms_per_step": msv, "GB/s": P.algo / msv / 1e6, "barrier_points_per_step": vl[0].n_barrier_points,
                        "note": "one cross-GPU barrier point per step only: nothing orders a matmul group behind the groups it would "
                                "consume in a model" if dep_chain else "every new activation vector waits for all earlier groups on all GPUs"}}
            del vl
        except Exception as e:
            variant = {"error": repr(e)}
        barrier()

    # ---- second leg for row split: the 70B shapes of BASELINE configs[4] (the north star's scaling target) -----
    leg70 = None
    if os.environ.get("B200Q_BENCH_70B", "1" if world > 1 else "0") == "1" and args.workload != "llama3-70b-q4_k":
        try:
            del lists, graphs, chunk_lists, e2e_graphs
            P70 = Pass("llama3-70b-q4_k", seed0=5678)
            l70 = [P70.make_list(b, dep=dep_chain) for b in range(2)]
            n70 = [0]

            def step70():
                l70[n70[0] & 1].launch()
                n70[0] += 1
            for _ in range(4):
                step70()
            barrier()
            k70 = max(4, min(args.steps, 40)) & ~1
            e0.record()
            for _ in range(k70):
                step70()
            e1.record()
            barrier()
            ms70 = reduce_max(e0.elapsed_time(e1)) / k70
            leg70 = {"workload": "llama3-70b-q4_k decode GEMV set, n=1 (561 matmuls, 39.4 GB)", "ms_per_step": ms70,
                     "value": P70.algo / ms70 / 1e6, "unit": UNIT, "steps": k70, "launches_per_step": l70[0].n_launches,
                     "barrier_points_per_step": l70[0].n_barrier_points, "per_gpu_frac_of_measured_hbm": P70.algo / ms70 / 1e6 / world / measured_peaks()[0]}
        except Exception as e:
            leg70 = {"error": repr(e)}
        barrier()

    def leave():
        """all ranks: flush, meet once more, then leave without tearing NCCL down (destroy_process_group can hang behind
        captured graphs that still reference the communicator)"""
        sys.stdout.flush()
        sys.stderr.flush()
        if dist is not None:
            torch.cuda.synchronize()
            dist.barrier()
            os._exit(0)
        return 0

    if rank != 0:
        return leave()

    algo = P.algo
    value = algo / (ms_dev * 1e-3) / 1e9
    e2e_value = algo / (ms_e2e * 1e-3) / 1e9
    peak, peak_src = measured_peaks()
    bf16_peak = measured_bf16()
    info = P.make_list(0, dep=dep_chain)
    n_launch = info.n_launches
    desc = workloads.describe(ops)
    chain_txt = ("every new activation vector waits for all earlier matmuls (model order: qkv -> wo -> gate/up -> down -> next layer), "
                 f"{info.n_barrier_points} device-side barrier points per step" if dep_chain else "no dependencies between matmuls (upper bound)")
    # DRAM bytes of ONE launch of the dominant kernel from the committed ncu --set full capture of the same kernel on the same
    # workload (profiles/list_traffic.json; ncu cannot run inside a timed bench): reported only for the configuration it was taken on
    traffic, traffic_note = None, "no ncu capture for this configuration"
    try:
        with open(os.path.join(ROOT, "profiles", "list_traffic.json")) as f:
            tj = json.load(f).get("full_step")
        if tj and world == 1 and n_launch == 1 and args.workload == tj.get("workload") and int(algo) == int(tj["algorithmic_bytes"]):
            traffic = tj["dram_bytes_read"] + tj["dram_bytes_write"]
            traffic_note = f"bytes per launch; {tj['source']}: dram read + write = {tj['ratio']} x algorithmic bytes (not measured in this run)"
        elif tj:
            traffic_note = "not measured for this configuration; single-GPU whole-step capture: " + tj["source"]
    except Exception as e:  # noqa: BLE001
        traffic_note = f"profiles/list_traffic.json unreadable: {e!r}"
    out = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms_dev, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
        "dtype": "int8 dp4a dot (q8 activations) / f32 accumulate", "data": "synthetic",
        "config": {"workload": args.workload + " decode GEMV set, n=1, tg (one step = the 225 matmuls of one token)",
                   "n_matmuls": desc["n_matmuls"], "weight_bytes": desc["weight_bytes"], "shapes": desc["shapes"],
                   "l2": "inputs larger than L2 (4.6 GB of distinct weights per step), no flush needed",
                   "dependencies": chain_txt,
                   "parallelism": "single GPU, one persistent kernel launch per step (b200q_list)" if world == 1 else (
                       f"row-split x{world}: one persistent launch per rank and step, result rows stored to all peers by the kernel "
                       f"(NVLink), consecutive matmul groups are DEPENDENT across GPUs through {info.n_barrier_points} per-group arrival "
                       "flags (no collective, no host sync)"),
                   "tg_tok_s_matmul_only": 1e3 / ms_dev},
        "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                "ms_per_step": ms_e2e, "results_finite": finite, "matches_resident_pass": e2e_ok,
                "api": f"b200q_memcpy(H2D) + {len(cuts) - 1} x b200q_list_launch ({len(ops)} matmuls) + b200q_memcpy(D2H) + "
                       "b200q_stream_synchronize",
                "mode": f"one CUDA graph per step: activations H2D on a copy stream (2 joins), {len(cuts) - 1} list launches, results "
                        "D2H on a second copy stream after each launch" + ("" if world == 1 else "; every rank copies all activations in, rank 0 copies the results out")},
        "gpu_launches": n_launch * args.steps,
        "roofline": {"bound": "hbm", "achieved": value / world, "peak": peak, "unit": "GB/s",
                     "frac": value / world / peak, "traffic": traffic, "peak_source": peak_src,
                     "traffic_note": traffic_note,
                     "kernel": "gemv_list_kernel (one launch = the whole step)",
                     "launches_per_step": n_launch,
                     "algorithmic_bytes_per_launch_avg": algo / world / max(1, n_launch),
                     "avg_launch_us": ms_dev * 1e3 / max(1, n_launch), "frac_of_nominal_8TBs": value / world / 8000.0},
        "clocks": clocks,
    }
    if rowsplit_parity is not None:
        out["rowsplit_parity"] = rowsplit_parity
    if leg70 is not None:
        out.setdefault("aux", {})["config5_llama3_70b_rowsplit"] = leg70
    if variant is not None:
        out.setdefault("aux", {})["step_variants"] = variant
    if world == 1:
        aux = out.setdefault("aux", {})
        try:
            # the same step without barrier points (upper bound) and as round 1's chain of grouped launches
            nd = P.make_list(0, dep=not dep_chain)
            ms_nd = _timeit(torch, nd.launch, 20)
            aux["step_variants"] = {("no_dependencies" if dep_chain else "with_dependencies"): {"ms_per_step": ms_nd, "GB/s": algo / ms_nd / 1e6}}

            def chain():
                i = 0
                while i < len(ops):
                    j = i + 1
                    while j < len(ops) and j - i < 4 and P.same_vec(j) and ops[j].qtype == ops[i].qtype:
                        j += 1
                    b200q.mul_mat_group([P.shards[q].shard for q in range(i, j)], P.act_dev[P.k_off[i]:P.k_off[i + 1]].view(1, -1),
                                        outs=[P.out[0][P.m_off[q]:P.m_off[q + 1]].view(1, -1) for q in range(i, j)])
                    i = j
            chain()
            torch.cuda.synchronize()
            gc = torch.cuda.CUDAGraph()
            with torch.cuda.graph(gc):
                chain()
            ms_c = _timeit(torch, gc.replay, 20)
            aux["step_variants"]["round1_chain_of_grouped_launches"] = {"ms_per_step": ms_c, "GB/s": algo / ms_c / 1e6}
        except Exception as e:
            aux["step_variants"] = {"error": repr(e)}
        try:
            aux.update(aux_single_gpu(torch, b200q, workloads, peak, bf16_peak))
        except Exception as e:  # never lose the headline line to the aux leg
            aux["error"] = repr(e)
    if world == 1 and not args.no_cpu_baseline:
        out["cpu_baseline"] = run_cpu_reference(cpu_sample_ops(ops), args.cpu_seconds)
    if world == 1 and not args.no_llama_bench:
        torch.cuda.empty_cache()
        out.setdefault("aux", {})["llama_bench_8b_q4_k_m"] = aux_llama_bench()
    if rowsplit_parity not in (None, "bit-exact"):
        out["error"] = "row-split result differs from the single-GPU result"
    sys.stdout.flush()
    os.dup2(saved_stdout, 1)
    print(json.dumps(out))
    if rowsplit_parity not in (None, "bit-exact"):
        sys.stdout.flush()
        os._exit(3)
    return leave()


if __name__ == "__main__":
    a = parse_args()
    sys.exit(main_reference(a) if a.impl == "reference" else main_b200(a))
