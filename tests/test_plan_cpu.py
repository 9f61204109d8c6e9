"""Host logic of the multi-column decode kernel (csrc/api.cu plan_mma), no GPU needed: every plan must respect the
invariants the kernel relies on (csrc/gemv_mma.cu) — ring slots for every warp, the bank-conflict-free pitches, k-split
teams that divide the chunks, 16-byte bulk-copy windows that fit the pitch, shared memory within the device limit."""
import ctypes as C

import pytest

import refs

SMS, MAX_SMEM = 148, 232448  # B200
NAMES = ["nwarps", "cluster", "kc", "chunks", "kp", "cpm", "pitch", "stage_bytes", "nstages", "groups", "grid", "smem",
         "col_pitch", "off_stage", "chunk_bytes", "row_bytes"]
TYPES = [refs.Q4_K, refs.Q5_K, refs.Q6_K, refs.Q4_0, refs.Q4_1, refs.Q5_0, refs.Q5_1, refs.Q8_0]
SHAPES = [(1, 256), (17, 768), (33, 1024), (100, 2048), (4096, 4096), (1024, 4096), (14336, 4096), (4096, 14336),
          (128256, 4096), (8192, 8192), (28672, 8192), (8192, 28672), (40, 28672)]


def plan(t, m, k, n, sms=SMS, max_smem=MAX_SMEM):
    import b200q
    fn = b200q.lib().b200q_debug_plan_gemv_mma
    fn.restype = C.c_int
    fn.argtypes = [C.c_int, C.c_int64, C.c_int64, C.c_int, C.c_int, C.c_int, C.POINTER(C.c_int)]
    out = (C.c_int * 16)()
    ok = fn(t, m, k, n, sms, max_smem, out)
    return dict(zip(NAMES, out)) if ok else None


@pytest.mark.parametrize("t", TYPES, ids=lambda t: refs.NAMES[t])
def test_plan_invariants(t):
    kquant = t in (refs.Q4_K, refs.Q5_K, refs.Q6_K)
    planned = 0
    for m, k in SHAPES:
        for n in (2, 3, 5, 8):
            p = plan(t, m, k, n)
            row_bytes = refs.row_size(t, k)
            if p is None:
                # rows that are not 16-byte multiples (or images that do not fit) fall back to the per-column kernel
                assert row_bytes % 16 != 0 or n * k > 150000, (m, k, n)
                continue
            planned += 1
            assert p["row_bytes"] == row_bytes
            assert p["nwarps"] in (8, 16) and p["nstages"] >= p["nwarps"], p
            assert p["smem"] <= MAX_SMEM and p["off_stage"] % 128 == 0, p
            assert p["off_stage"] + p["nstages"] * p["stage_bytes"] <= p["smem"], p
            unit = 256 if kquant else 32
            assert p["kc"] * p["chunks"] == k // unit and p["chunk_bytes"] * p["chunks"] == row_bytes, p
            assert p["chunks"] % p["kp"] == 0 and p["cpm"] * p["kp"] == p["chunks"], p
            assert p["kp"] in (1, 2, 4, 8) and p["nwarps"] % p["kp"] == 0, p
            assert p["kp"] == 1 or p["nwarps"] // p["kp"] <= 8, p   # named barriers 1..8
            # the 16-byte window around a chunk (lead <= 14 bytes) fits the pitch; the pitch keeps fragment loads conflict-free
            assert p["pitch"] % 16 == 0 and p["pitch"] >= (p["chunk_bytes"] + 14 + 15) // 16 * 16, p
            assert p["stage_bytes"] == 16 * p["pitch"], p
            if t in (refs.Q4_K, refs.Q5_K):
                assert p["pitch"] % 128 in (32, 96), p
            else:
                assert (p["pitch"] // 16) % 2 == 1, p
            assert p["col_pitch"] == k + (64 if kquant else 16), p
            assert p["groups"] == (m + 15) // 16 and 1 <= p["grid"] <= min(p["groups"], SMS), p
            assert p["cluster"] in (1, 2) and (p["cluster"] == 1 or p["grid"] % 2 == 0), p
    assert planned > 0


def test_plan_rejects_bad_shapes():
    assert plan(refs.Q4_K, 4096, 4096, 1) is None       # n = 1 belongs to gemv_kq
    assert plan(refs.Q4_K, 4096, 4096, 9) is None       # more than 8 columns: the caller splits
    assert plan(refs.Q4_K, 4096, 4000, 4) is None       # k not a multiple of the super-block
    assert plan(refs.MXFP4, 4096, 4096, 4) is None      # no tile for this format
    assert plan(refs.Q4_K, 64, 28672, 8) is None        # eight int8 images of 28 KB do not fit beside a ring
    assert plan(refs.Q4_K, 64, 28672, 4) is not None    # ... four do (the caller halves the columns)


def test_plan_uses_the_machine():
    """long matrices keep 16 warps with k-split teams only when that shortens the per-warp chain; short matrices split k"""
    head = plan(refs.Q4_K, 128256, 4096, 8)
    assert head["nwarps"] == 16 and head["grid"] == SMS and head["cluster"] == 2
    wo = plan(refs.Q4_K, 4096, 4096, 8)
    assert wo["kp"] >= 4                                  # 256 row groups over 148 CTAs: k is split across warps
    down = plan(refs.Q4_K, 4096, 14336, 2)
    assert down["kc"] == 2 and down["kp"] == 4            # half-size stages so that four warps share a group


def _workspace(t, m, k, n, flags=0, batch=1):
    import b200q
    rs = refs.row_size(t, k)
    a = b200q.MulMatArgs(t, k, m, n, 1, 1, batch, 1, rs, rs * m, rs * m, k * 4, k * 4 * n, k * 4 * n * batch,
                         m * 4, m * 4 * n, m * 4 * n * batch, flags)
    fn = b200q.lib().b200q_mul_mat_workspace
    fn.restype = C.c_size_t
    return fn(C.byref(a))


def test_kernel_choice_by_columns():
    """b200q_mul_mat_workspace mirrors the dispatch (api.cu use_mmq): 0 bytes = a GEMV kernel serves the shape, > 0 = the
    tensor-core kernel (which needs its packed bf16 activation tiles).  Pure host logic."""
    import b200q
    m, k = 4096, 4096
    for n in (1, 2, 8, 9, 16):
        assert _workspace(refs.Q4_K, m, k, n) == 0, n            # gemv_kq / gemv_mma (two passes for 9..16)
    assert _workspace(refs.Q4_K, m, k, 17) > 0                   # measured cross-over: three passes no longer win
    assert _workspace(refs.Q4_K, m, k, 512) > 0
    assert _workspace(refs.MXFP4, m, k, 8) == 0 and _workspace(refs.MXFP4, m, k, 9) > 0      # no MMA GEMV tile
    assert _workspace(refs.Q6_K, m, 1024, 8) == 0 and _workspace(refs.Q6_K, m, 1024, 9) > 0  # 840-byte rows: no plan
    assert _workspace(refs.Q4_K, m, k, 12, batch=3) > 0          # batch dims: no multi-pass GEMV
    assert _workspace(refs.Q4_K, m, k, 4, flags=b200q.FLAG_FORCE_MMQ) > 0
    assert _workspace(refs.Q4_K, m, k, 512, flags=b200q.FLAG_FORCE_GEMV) == 0


# ---- the persistent decode list (csrc/api.cu list_plan / gemv_list.cu) ------------------------------------------------
LIST_PAYLOAD, LIST_MIN_SLOTS = 9216, 16  # csrc/gemv_list.cuh (8 warps)


def plan_list(ops, world=1, sms=SMS, max_smem=MAX_SMEM):
    """ops: [(type, m, k, dep, same_act_as_previous)] -> (per-op dicts, per-launch dicts)"""
    import b200q
    fn = b200q.lib().b200q_debug_plan_list
    fn.restype = C.c_int
    n = len(ops)
    ty = (C.c_int * n)(*[o[0] for o in ops])
    mm = (C.c_int64 * n)(*[o[1] for o in ops])
    kk = (C.c_int64 * n)(*[o[2] for o in ops])
    dep = (C.c_int * n)(*[int(o[3]) for o in ops])
    same = (C.c_int * n)(*[int(o[4]) for o in ops])
    out = (C.c_int * (8 * n))()
    seg = (C.c_int * (8 * 64))()
    rc = fn(ty, mm, kk, dep, same, n, world, sms, max_smem, out, seg, 64)
    assert rc >= 1, (rc, b200q.lib().b200q_last_error())
    on = ["launch", "units", "rps", "nch", "csb", "gen", "nsb", "rot"]
    sn = ["n_ops", "n_gens", "total_units", "n_bp", "nslots", "smem", "grid", "act_stride"]
    return ([dict(zip(on, out[8 * i:8 * i + 8])) for i in range(n)], [dict(zip(sn, seg[8 * i:8 * i + 8])) for i in range(rc)])


def _model_ops(workload):
    from b200q import workloads
    ops = []
    for o in workload:
        leaf = o.name.split(".")[-1]
        same = leaf in ("attn_k", "attn_v", "ffn_up")
        ops.append((o.qtype, o.m, o.k, not same, same))
    return ops


def test_list_plan_llama3_8b_is_one_launch():
    from b200q import workloads
    ops = _model_ops(workloads.llama3_8b("q4_k_m"))
    per_op, segs = plan_list(ops)
    assert len(segs) == 1, segs
    s = segs[0]
    assert s["n_ops"] == 225 and s["n_gens"] == 129 and s["n_bp"] == 128  # the first generation follows a kernel boundary
    assert s["nslots"] >= LIST_MIN_SLOTS and s["smem"] <= MAX_SMEM and s["grid"] == SMS
    units = 0
    for (t, m, k, _, _), p in zip(ops, per_op):
        rb = refs.row_size(t, k)
        assert p["units"] == (m + p["rps"] - 1) // p["rps"] and 0 <= p["rot"] < s["grid"]
        if p["nch"] == 1:
            assert p["rps"] * rb <= LIST_PAYLOAD and p["csb"] == p["nsb"]
        else:
            assert p["rps"] == 1 and p["csb"] * refs.BLCK[t][1] <= LIST_PAYLOAD and p["csb"] * p["nch"] >= p["nsb"] > p["csb"] * (p["nch"] - 1)
        units += (m + p["rps"] - 1) // p["rps"]
    assert units == s["total_units"]


def test_list_plan_llama3_70b_splits_and_chunks():
    from b200q import workloads
    ops = _model_ops(workloads.llama3_70b("q4_k"))
    per_op, segs = plan_list(ops, world=8)
    assert 1 <= len(segs) <= 4
    assert sum(s["n_ops"] for s in segs) == len(ops)
    for s in segs:
        assert s["nslots"] >= LIST_MIN_SLOTS and s["smem"] <= MAX_SMEM and s["n_bp"] >= 1  # row split: a final barrier point per launch
    down = [p for (t, m, k, _, _), p in zip(ops, per_op) if k == 28672]
    assert down and all(p["nch"] == 2 and p["rps"] == 1 for p in down)


@pytest.mark.parametrize("t", [refs.Q4_K, refs.Q5_K, refs.Q6_K, refs.Q8_0, refs.Q4_0, refs.Q4_1, refs.Q5_0, refs.Q5_1, refs.MXFP4],
                         ids=lambda t: refs.NAMES[t])
def test_list_geometry_depends_on_type_and_k_only(t):
    """the same (type, k) gets the same stage geometry whatever the list looks like (results must not depend on it)"""
    unit = 256 if t in (refs.Q4_K, refs.Q5_K, refs.Q6_K) else 32
    for k in (unit, 4 * unit, 2880 // 32 * 32 if unit == 32 else 1024, 4096, 14336 // unit * unit, 28672):
        if k % unit:
            continue
        a, _ = plan_list([(t, 77, k, False, False)])
        b, _ = plan_list([(refs.Q4_K, 4096, 4096, False, False), (t, 5000, k, True, False), (refs.Q6_K, 128, 14336, False, False)])
        for f in ("rps", "nch", "csb", "nsb"):
            assert a[0][f] == b[1][f], (k, f)


def test_list_q5_k_rows_are_walked_by_lanes_not_columns():
    """q5_K keeps its activations in shared memory (gemv_list.cu: the register-resident column scheme is q4_K / q6_K only), so a
    stage must never hold more than 32 rows — one lane per row at most — however short the rows are"""
    for k in (256, 512, 1024, 4096, 14336):
        p, segs = plan_list([(refs.Q5_K, 300, k, False, False)])
        assert 1 <= p[0]["rps"] <= 32 and p[0]["rps"] * refs.row_size(refs.Q5_K, k) <= LIST_PAYLOAD or p[0]["nch"] > 1, (k, p[0])
        assert p[0]["nsb"] == k // 256 and segs[0]["nslots"] >= LIST_MIN_SLOTS


def test_list_walk_covers_every_unit_exactly_once():
    """emulation of gemv_list.cu's per-warp walk (next_op / advance) over the planned table: every unit of every op is
    taken by exactly one (CTA, warp), the CTAs' loads differ by at most one unit over the whole launch"""
    import numpy as np
    from b200q import workloads
    ops = _model_ops(workloads.llama3_8b("q4_k_m", layers=3)) + [(refs.Q4_0, 5, 1024, False, False), (refs.Q8_0, 0, 1024, False, False),
                                                                 (refs.Q6_K, 1000, 2048, True, False)]
    per_op, segs = plan_list(ops)
    assert len(segs) == 1
    G, NW = segs[0]["grid"], 8
    seen = [np.zeros(p["units"], dtype=np.int32) for p in per_op]
    load = np.zeros(G, dtype=np.int64)
    for c in range(G):
        for w in range(NW):
            woff = 0
            for j, p in enumerate(per_op):
                base, rem = p["units"] // G, p["units"] % G
                cp = c + p["rot"]
                cp -= G if cp >= G else 0
                ncta = base + (1 if cp < rem else 0)
                first = (w - woff) & (NW - 1)
                woff = (woff + ncta) & (NW - 1)
                if ncta > first:
                    start = cp * base + min(cp, rem) + first
                    left = (ncta - first + NW - 1) // NW
                    for i in range(left):
                        seen[j][start + i * NW] += 1
                    load[c] += left * p["nch"]
    for j, sarr in enumerate(seen):
        assert (sarr == 1).all(), j
    nch_max = max(p["nch"] for p in per_op)
    assert load.max() - load.min() <= 2 * nch_max + 1, (load.min(), load.max())  # units are balanced to 1; a chunked unit is nch stages
