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
