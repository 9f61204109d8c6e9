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
    ops = workload_ops(args.workload)

    # ---- resident state: weights (row shard of every matrix when world > 1), activations, results ------------
    from b200q import rowsplit
    shards = []
    for i, o in enumerate(ops):
        parts = rowsplit.row_split(o.m, world)
        row0, nrows = parts[rank]
        shard = workloads.synth_weights_device(o, seed=1234 + i, rows=nrows, row0=row0)
        shards.append(rowsplit.RowSplitWeight(o.qtype, o.m, o.k, rank, world, parts, shard))
    k_off, m_off = [0], [0]
    for o in ops:
        k_off.append(k_off[-1] + o.k)
        m_off.append(m_off[-1] + o.m)
    g = torch.Generator(device="cuda")
    g.manual_seed(4321)
    act_dev = torch.rand(k_off[-1], device="cuda", generator=g) * 2 - 1          # all activation vectors, f32
    # row split (N > 1): "peer" = the decode kernel stores its rows into every rank's result buffer itself (NVLink peer
    # stores in the epilogue, one small all-reduce per step as the completion fence); "nccl" = all-gather per matmul
    split_mode = os.environ.get("B200Q_ROWSPLIT", "peer") if world > 1 else "single"
    peer_dst = None
    if split_mode == "peer":
        peer_dst = rowsplit.PeerDst(m_off[-1], rank, world)
        out_dev = peer_dst.tensor
        out_dev.zero_()
    else:
        out_dev = torch.zeros(m_off[-1], device="cuda")                          # all (gathered) results, f32
    act_host = act_dev.cpu().pin_memory()
    out_host = torch.empty(m_off[-1], dtype=torch.float32).pin_memory()

    # launches of one pass: matrices of one type that multiply the same activation vector in the model (wq/wk/wv of a
    # layer, ffn_gate/ffn_up) go out as ONE grouped launch (b200q_mul_mat_group) and read the same vector here too
    SHARED = ({"attn_q", "attn_k", "attn_v"}, {"ffn_gate", "ffn_up"})
    launches = []
    for i, o in enumerate(ops):
        leaf = o.name.split(".")[-1]
        prev = launches[-1] if launches else None
        if (split_mode != "nccl" and prev is not None and len(prev) < 4 and ops[prev[0]].qtype == o.qtype and ops[prev[0]].k == o.k
                and ops[prev[0]].name.rsplit(".", 1)[0] == o.name.rsplit(".", 1)[0]
                and any(leaf in grp and ops[prev[0]].name.split(".")[-1] in grp for grp in SHARED)):
            prev.append(i)
        else:
            launches.append([i])

    def run_pass():
        """one pass over the matmul list; every launch goes to torch's current stream through the C-ABI"""
        for idx in launches:
            i = idx[0]
            a = act_dev[k_off[i]:k_off[i + 1]].view(1, ops[i].k)
            if world == 1:
                b200q.mul_mat_group([shards[j].shard for j in idx], a,
                                    outs=[out_dev[m_off[j]:m_off[j + 1]].view(1, ops[j].m) for j in idx])
            elif split_mode == "peer":
                rowsplit.mul_mat_row_split_peer([shards[j] for j in idx], a, peer_dst, [m_off[j] for j in idx])
            else:
                # this rank's rows + NCCL all-gather of the slices (rowsplit.py)
                rowsplit.mul_mat_row_split(shards[i], a, out=out_dev[m_off[i]:m_off[i + 1]].view(1, ops[i].m))
        if peer_dst is not None:
            peer_dst.fence()  # every rank's rows have landed in every rank's buffer

    for _ in range(2):
        run_pass()
    torch.cuda.synchronize()
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph):
        run_pass()
    torch.cuda.synchronize()

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

    # ---- value: device-resident ---------------------------------------------------------------------------
    for _ in range(max(3, args.warmup)):
        graph.replay()
    barrier()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.steps):
        graph.replay()
    e1.record()
    barrier()
    ms_dev = reduce_max(e0.elapsed_time(e1)) / args.steps
    clocks = sampler.stop() if rank == 0 else None

    # ---- e2e: host buffers through the C-ABI, copies inside the timed region --------------------------------
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    h2d = act_host.numel() * 4
    d2h = out_host.numel() * 4

    def e2e_step_serial():
        rc = L.b200q_memcpy(C.c_void_p(act_dev.data_ptr()), C.c_void_p(act_host.data_ptr()), C.c_size_t(h2d), 1, stream)
        assert rc == 0
        graph.replay()
        rc = L.b200q_memcpy(C.c_void_p(out_host.data_ptr()), C.c_void_p(out_dev.data_ptr()), C.c_size_t(d2h), 2, stream)
        assert rc == 0
        assert L.b200q_stream_synchronize(stream) == 0

    # single GPU: the same copies and launches as ONE CUDA graph in which the copies ride two copy streams — the
    # activations of launch chunk c+1 arrive while chunk c computes, the results of chunk c leave while chunk c+1
    # computes (every step still moves all h2d + d2h bytes inside the timed region)
    e2e_mode, e2e_graph = "serial", None
    if world == 1 and os.environ.get("B200Q_BENCH_E2E", "pipelined") == "pipelined":
        try:
            n_chunks = int(os.environ.get("B200Q_BENCH_E2E_CHUNKS", "16"))
            per = (len(launches) + n_chunks - 1) // n_chunks
            chunks = [launches[i:i + per] for i in range(0, len(launches), per)]
            s_in, s_out = torch.cuda.Stream(), torch.cuda.Stream()

            def copy(dst_t, src_t, lo, hi, kind, st):
                rc = L.b200q_memcpy(C.c_void_p(dst_t.data_ptr() + 4 * lo), C.c_void_p(src_t.data_ptr() + 4 * lo),
                                    C.c_size_t(4 * (hi - lo)), kind, C.c_void_p(st.cuda_stream))
                assert rc == 0

            e2e_graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(e2e_graph):
                main = torch.cuda.current_stream()
                s_in.wait_stream(main)
                s_out.wait_stream(main)
                # activations: the first chunk's slice, then all the rest — only TWO joins into the compute stream (every
                # join costs the programmatic-launch overlap of the kernel behind it); results leave chunk by chunk
                n_join = int(os.environ.get("B200Q_BENCH_E2E_JOINS", "2"))
                first = [j for idx in chunks[0] for j in idx]
                arrived = {}
                if n_join <= 2:
                    copy(act_dev, act_host, k_off[first[0]], k_off[first[-1] + 1], 1, s_in)
                    arrived[0] = torch.cuda.Event()
                    arrived[0].record(s_in)
                    if len(chunks) > 1:
                        copy(act_dev, act_host, k_off[first[-1] + 1], k_off[-1], 1, s_in)
                        arrived[1] = torch.cuda.Event()
                        arrived[1].record(s_in)
                else:
                    for c, ch in enumerate(chunks):
                        ops_c = [j for idx in ch for j in idx]
                        copy(act_dev, act_host, k_off[ops_c[0]], k_off[ops_c[-1] + 1], 1, s_in)
                        arrived[c] = torch.cuda.Event()
                        arrived[c].record(s_in)
                for c, ch in enumerate(chunks):
                    ops_c = [j for idx in ch for j in idx]
                    if c in arrived:
                        main.wait_event(arrived[c])
                    for idx in ch:
                        i = idx[0]
                        b200q.mul_mat_group([shards[j].shard for j in idx], act_dev[k_off[i]:k_off[i + 1]].view(1, ops[i].k),
                                            outs=[out_dev[m_off[j]:m_off[j + 1]].view(1, ops[j].m) for j in idx])
                    done = torch.cuda.Event()
                    done.record(main)
                    s_out.wait_event(done)
                    copy(out_host, out_dev, m_off[ops_c[0]], m_off[ops_c[-1] + 1], 2, s_out)
                main.wait_stream(s_in)
                main.wait_stream(s_out)
            # the pipelined graph must move exactly what the serial step moves
            out_host.zero_()
            e2e_graph.replay()
            torch.cuda.synchronize()
            chk = out_host.clone()
            e2e_step_serial()
            assert torch.equal(chk, out_host), "pipelined e2e graph and serial e2e step disagree"
            e2e_mode = (f"pipelined (one CUDA graph: activations H2D on a copy stream with {len(arrived)} joins, results D2H on a "
                        f"second copy stream after each of {len(chunks)} chunks of launches)")
        except Exception as e:  # keep the serial step rather than lose the bench line
            print(f"[bench] pipelined e2e unavailable ({e!r}); using the serial step", file=sys.stderr)
            e2e_graph = None
            torch.cuda.synchronize()

    def e2e_step():
        if e2e_graph is None:
            return e2e_step_serial()
        e2e_graph.replay()
        assert L.b200q_stream_synchronize(stream) == 0

    for _ in range(max(3, args.warmup)):
        e2e_step()
    barrier()
    t0 = time.perf_counter()
    e0.record()
    for _ in range(args.steps):
        e2e_step()
    e1.record()
    barrier()
    ms_e2e = reduce_max(max(e0.elapsed_time(e1), (time.perf_counter() - t0) * 1e3 * 0.0)) / args.steps
    finite = bool(torch.isfinite(out_host).all())

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

    algo = sum(o.algo_bytes(1) for o in ops)
    value = algo / (ms_dev * 1e-3) / 1e9
    e2e_value = algo / (ms_e2e * 1e-3) / 1e9
    peak, peak_src = measured_peaks()
    n_launch = len(launches)
    desc = workloads.describe(ops)
    traffic = None
    tpath = os.path.join(ROOT, "profiles", "gemv_traffic.json")
    if os.path.exists(tpath):
        with open(tpath) as f:
            tj = json.load(f)
            # ncu --set full capture of 8 consecutive launches (one layer): dram bytes / algorithmic bytes, applied to the
            # step's average launch
            traffic = tj.get("dram_over_algorithmic", 0) * algo / n_launch or None
    out = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms_dev, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
        "dtype": "int8 dp4a dot (q8 activations) / f32 accumulate", "data": "synthetic",
        "config": {"workload": args.workload + " decode GEMV set, n=1, tg (one step = the 225 matmuls of one token)",
                   "n_matmuls": desc["n_matmuls"], "weight_bytes": desc["weight_bytes"], "shapes": desc["shapes"],
                   "l2": "inputs larger than L2 (4.6 GB of distinct weights per step), no flush needed",
                   "parallelism": "single GPU" if world == 1 else (
                       f"row-split x{world}, result rows stored to all peers by the kernel (NVLink), one all-reduce fence per step"
                       if split_mode == "peer" else f"row-split x{world} + NCCL all-gather per matmul"),
                   "tg_tok_s_matmul_only": 1e3 / ms_dev},
        "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                "ms_per_step": ms_e2e, "results_finite": finite,
                "api": f"b200q_memcpy(H2D) + {n_launch} x b200q_mul_mat_group ({len(ops)} matmuls) + b200q_memcpy(D2H) + "
                       "b200q_stream_synchronize", "mode": e2e_mode},
        "gpu_launches": n_launch * args.steps,
        "roofline": {"bound": "hbm", "achieved": value / world, "peak": peak, "unit": "GB/s",
                     "frac": value / world / peak, "traffic": traffic, "peak_source": peak_src,
                     "kernel": f"gemv_kq_kernel<q4_K | q6_K> (all {n_launch} launches of the step; wq/wk/wv and gate/up grouped)",
                     "launches_per_step": n_launch,
                     "algorithmic_bytes_per_launch_avg": algo / n_launch,
                     "avg_launch_us": ms_dev * 1e3 / n_launch, "frac_of_nominal_8TBs": value / world / 8000.0},
        "clocks": clocks,
    }
    if world == 1:
        # aux (not the headline): the longest launch of the step alone — the output head, 430 MB > L2, so every launch
        # streams from HBM — on the n = 1 kernel and, with 8 activation columns, on the MMA tile kernel (gemv_mma.cu)
        try:
            head = max(range(len(ops)), key=lambda i: ops[i].m * ops[i].k)
            o, wq = ops[head], shards[head].shard
            aux = []
            for n, kern in ((1, "gemv_kq_kernel"), (8, "gemv_mma_kernel")):
                a = (torch.rand(n, o.k, device="cuda") * 2 - 1)
                dst = torch.empty(n, o.m, device="cuda")
                for _ in range(3):
                    b200q.mul_mat(wq, a, out=dst)
                torch.cuda.synchronize()
                e0.record()
                for _ in range(20):
                    b200q.mul_mat(wq, a, out=dst)
                e1.record()
                torch.cuda.synchronize()
                us = e0.elapsed_time(e1) * 1e3 / 20
                gbs = o.algo_bytes(n) / us / 1e3
                aux.append({"kernel": f"{kern}<{b200q.TYPE_NAMES.get(o.qtype, o.qtype)}>", "shape": f"{o.k}->{o.m}", "n": n, "us": us,
                            "achieved": gbs, "unit": "GB/s", "frac": gbs / peak})
            out["aux"] = {"single_launch": aux}
        except Exception as e:  # never lose the headline line to the aux leg
            out["aux"] = {"error": repr(e)}
    if world == 1 and not args.no_cpu_baseline:
        out["cpu_baseline"] = run_cpu_reference(cpu_sample_ops(ops), args.cpu_seconds)
    sys.stdout.flush()
    os.dup2(saved_stdout, 1)
    print(json.dumps(out))
    return leave()


if __name__ == "__main__":
    a = parse_args()
    sys.exit(main_reference(a) if a.impl == "reference" else main_b200(a))
