"""Pins the oracle (oracle/oracle.c) to the reference:
   * against the committed golden vectors (tests/golden/ref_vectors.npz, produced by the unmodified reference
     through tests/golden/make_golden.py) — runs everywhere;
   * live against oracle/_ref (the reference compiled here) when it is present.
Bit-exact for dequantize / GET_ROWS / activation quantizers (fp32 d*q arithmetic); fp32-rounding
tolerance for dot products (the reference's SIMD kernels sum in a different order than its *_generic ones)."""
import os

import numpy as np
import pytest

import refs

GOLD = np.load(os.path.join(os.path.dirname(__file__), "golden", "ref_vectors.npz"))
M, K, N, E = 8, 512, 3, 4


@pytest.mark.parametrize("t", refs.WEIGHT_TYPES, ids=lambda t: refs.NAMES[t])
def test_dequantize_bit_exact_vs_golden(t):
    name = refs.NAMES[t]
    for tag in ("", "_cos"):
        got = refs.orc_dequantize(t, GOLD[f"{name}/wq{tag}"], K)
        want = GOLD[f"{name}/deq{tag}"]
        assert got.tobytes() == want.tobytes(), f"{name}{tag}: dequantize differs from the reference"


@pytest.mark.parametrize("t", refs.WEIGHT_TYPES, ids=lambda t: refs.NAMES[t])
def test_activation_quantizer_vs_golden(t):
    """quantize_row_q8_0/q8_1/q8_K: the x86 SIMD quantizers round half-to-even where *_ref uses roundf, and q8_1
    stores d*sum in fp16 — allow one code of difference on <0.1 % of the values, scales bit-exact."""
    name = refs.NAMES[t]
    vt = int(GOLD[f"{name}/vec_dot_type"][0])
    act = GOLD[f"{name}/act"]
    want = GOLD[f"{name}/actq"]
    got = np.zeros_like(want)
    for j in range(N):
        assert refs.oracle().orc_quantize_row(vt, refs._p(np.ascontiguousarray(act[j])), refs._p(got[j]), K) == 0
    be, bb = refs.BLCK[vt]
    g = got.reshape(N, K // be, bb)
    w = want.reshape(N, K // be, bb)
    hdr = {refs.Q8_0: 2, refs.Q8_1: 2, refs.Q8_K: 4}[vt]
    qoff = {refs.Q8_0: 2, refs.Q8_1: 4, refs.Q8_K: 4}[vt]
    assert np.array_equal(g[:, :, :hdr], w[:, :, :hdr]), "scale bytes differ"
    gq = g[:, :, qoff:qoff + be].view(np.int8).astype(np.int32)
    wq = w[:, :, qoff:qoff + be].view(np.int8).astype(np.int32)
    diff = np.abs(gq - wq)
    assert diff.max() <= 1 and (diff > 0).mean() < 1e-3


@pytest.mark.parametrize("t", refs.WEIGHT_TYPES, ids=lambda t: refs.NAMES[t])
def test_vec_dot_and_mul_mat_vs_golden(t):
    name = refs.NAMES[t]
    wq, act, aq = GOLD[f"{name}/wq"], GOLD[f"{name}/act"], GOLD[f"{name}/actq"]
    dots = np.zeros((N, M), dtype=np.float32)
    for j in range(N):
        for i in range(M):
            o = np.zeros(1, dtype=np.float32)
            assert refs.oracle().orc_vec_dot(t, K, refs._p(np.ascontiguousarray(wq[i])),
                                             refs._p(np.ascontiguousarray(aq[j])), refs._p(o)) == 0
            dots[j, i] = o[0]
    # same quantized inputs on both sides: only the fp32 summation order differs
    np.testing.assert_allclose(dots, GOLD[f"{name}/dots"], rtol=2e-5, atol=2e-5)
    mm = refs.orc_mul_mat(t, wq, act)
    assert refs.nmse(mm, GOLD[f"{name}/mm"]) < 1e-6
    np.testing.assert_allclose(mm, GOLD[f"{name}/mm"], rtol=1e-2, atol=2e-2)


@pytest.mark.parametrize("t", refs.WEIGHT_TYPES, ids=lambda t: refs.NAMES[t])
def test_mul_mat_id_vs_golden(t):
    name = refs.NAMES[t]
    wq = GOLD[f"{name}/wq"].reshape(E, M // E, -1)
    got = refs.orc_mul_mat_id(t, wq, GOLD[f"{name}/b"], GOLD[f"{name}/ids"], 2)
    assert refs.nmse(got, GOLD[f"{name}/mmid"]) < 1e-6


@pytest.mark.parametrize("t", [refs.Q4_0, refs.Q8_0, refs.MXFP4], ids=lambda t: refs.NAMES[t])
def test_weight_quantizers_vs_golden(t):
    """quantize_row_{q4_0,q8_0,mxfp4}_ref restated: identical bytes on the reference's cos fixture and U(-1,1)."""
    name = refs.NAMES[t]
    for tag, src in (("", refs.uniform_weights(M, K, 1234 + t)),
                     ("_cos", (0.1 + 2.0 * np.cos(np.arange(M * K, dtype=np.float32) + np.float32(t))).astype(np.float32).reshape(M, K))):
        out = np.zeros((M, refs.row_size(t, K)), dtype=np.uint8)
        for i in range(M):
            assert refs.oracle().orc_quantize_row(t, refs._p(np.ascontiguousarray(src[i])), refs._p(out[i]), K) == 0
        assert np.array_equal(out, GOLD[f"{name}/wq{tag}"]), f"{name}{tag}"


def test_fp16_conversions_exhaustive():
    O = refs.oracle()
    h = np.arange(65536, dtype=np.uint16)
    f = h.view(np.float16).astype(np.float32)
    for i in range(0, 65536, 7):
        v = O.orc_fp16_to_fp32(int(h[i]))
        assert (np.isnan(v) and np.isnan(f[i])) or np.float32(v).tobytes() == f[i].tobytes()
    rng = np.random.default_rng(0)
    x = np.concatenate([rng.normal(size=4000).astype(np.float32) * np.float32(10.0) ** rng.integers(-8, 5, 4000),
                        np.array([0.0, -0.0, 65504.0, 65519.9, 65520.0, 1e-8, 5.96e-8, 2.98e-8, 2.99e-8, 6.1e-5],
                                 dtype=np.float32)]).astype(np.float32)
    want = x.astype(np.float16).view(np.uint16)
    for xi, wi in zip(x, want):
        assert O.orc_fp32_to_fp16(float(xi)) == int(wi), (xi, wi)


# ---------------------------------------------------------------------------------------------------------
# live checks against the reference build (skipped when oracle/_ref did not travel)
# ---------------------------------------------------------------------------------------------------------
needs_ref = pytest.mark.skipif(refs.ref() is None, reason="oracle/_ref not built on this machine")


@needs_ref
@pytest.mark.parametrize("t", refs.WEIGHT_TYPES, ids=lambda t: refs.NAMES[t])
def test_live_dequantize_random_blocks(t):
    """random payload bytes exercise every code point, negative scales and all K-quant scale packings"""
    k = 1024
    rows = refs.random_blocks(t, 16, k, seed=77 + t)
    want = np.empty((16, k), dtype=np.float32)
    assert refs.ref().ref_dequantize(t, refs._p(rows), refs._p(want), 16 * k) == 0
    got = refs.orc_dequantize(t, rows, k)
    assert got.tobytes() == want.tobytes()


@needs_ref
@pytest.mark.parametrize("t", refs.WEIGHT_TYPES, ids=lambda t: refs.NAMES[t])
@pytest.mark.parametrize("shape", [(16, 256, 1), (33, 1024, 5), (64, 2048, 2)])
def test_live_mul_mat(t, shape):
    m, k, n = shape
    w = refs.make_weights(t, m, k, seed=m + k + t)
    act = refs.uniform_weights(n, k, seed=5 + t)
    want, _ = refs.ref_mul_mat(t, w, act, threads=2)
    got = refs.orc_mul_mat(t, w, act)
    assert refs.nmse(got, want) < 1e-6


@needs_ref
def test_live_get_rows():
    t, k, m = refs.Q4_K, 512, 20
    w = refs.make_weights(t, m, k, seed=3)
    rows = np.array([3, 3, 19, 0, 7], dtype=np.int32)
    want = np.empty((5, k), dtype=np.float32)
    assert refs.ref().ref_get_rows(t, refs._p(w), refs._p(rows), refs._p(want), k, m, 5, 2) >= 0
    got = np.empty((5, k), dtype=np.float32)
    assert refs.oracle().orc_get_rows(t, refs._p(w), refs._p(rows), refs._p(got), k, 5) == 0
    assert got.tobytes() == want.tobytes()
