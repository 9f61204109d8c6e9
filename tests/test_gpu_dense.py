"""GPU parity of MUL_MAT with unquantized weights (f16 / bf16 / f32) on the shapes the tensor-core kernel does not take
(csrc/dense_mm.cu): any k, few columns, strided rows, broadcast batches — the reference's cases of
test-backend-ops.cpp:5750-5790 (k = 4, 67, 129, 193, 1056, 1057; views; permutations).  The checker is a float64 product of the
exactly converted weights (the reference's CPU path rounds the activations to f16 first, ggml-cpu.c vec_dot_type of f16, so
its own gate NMSE <= 5e-4 applies; the kernel keeps f32 activations and is held to 1e-6 here)."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
MAX_NMSE = 1e-6


def _weights(kind, m, k, seed):
    """-> (raw bytes [m, k * es] uint8, exact float64 values [m, k])"""
    rng = np.random.default_rng(seed)
    x = rng.uniform(-1, 1, size=(m, k)).astype(np.float32)
    if kind == "f32":
        return x.view(np.uint8).reshape(m, -1), x.astype(np.float64)
    if kind == "f16":
        h = x.astype(np.float16)
        return h.view(np.uint8).reshape(m, -1), h.astype(np.float64)
    u = (x.view(np.uint32) >> 16).astype(np.uint16)  # bf16: truncation is a valid bf16 value
    return u.view(np.uint8).reshape(m, -1), (u.astype(np.uint32) << 16).view(np.float32).astype(np.float64)


def _code(kind):
    import b200q
    return {"f32": b200q.F32, "f16": b200q.F16, "bf16": b200q.BF16}[kind]


def _nmse(got, want):
    d = got.astype(np.float64) - want
    return float((d * d).sum() / max((want * want).sum(), 1e-30))


SHAPES = [(16, 1, 4), (16, 16, 4), (16, 8, 4), (1056, 1, 193), (1056, 1, 67), (1057, 1, 129), (129, 1, 1057), (128, 1, 1056),
          (67, 9, 33), (64, 3, 64), (64, 8, 256), (200, 40, 100), (130, 70, 1), (1, 1, 1), (4096, 1, 128), (33, 5, 4096)]


@pytest.mark.parametrize("kind", ["f16", "bf16", "f32"])
@pytest.mark.parametrize("m,n,k", SHAPES)
def test_dense_mul_mat_any_k(kind, m, n, k):
    import torch
    import b200q
    raw, wv = _weights(kind, m, k, seed=m * 7 + n * 3 + k)
    act = np.random.default_rng(99 + k).uniform(-1, 1, size=(n, k)).astype(np.float32)
    got = b200q.mul_mat(b200q.upload(_code(kind), (m, k), raw), torch.from_numpy(act).cuda()).cpu().numpy()
    want = act.astype(np.float64) @ wv.T
    assert got.shape == (n, m) and np.isfinite(got).all()
    assert _nmse(got, want) <= MAX_NMSE


@pytest.mark.parametrize("kind", ["f16", "bf16", "f32"])
@pytest.mark.parametrize("n", [1, 5, 19])
def test_dense_mul_mat_broadcast_batches(kind, n):
    """weights [ne03 = 2, ne02 = 3], activations [ne13 = 4, ne12 = 6]: dst batch (i3, i2) reads weight batch (i3 / 2, i2 / 2)"""
    import torch
    import b200q
    m, k = 37, 75
    raw, wv = _weights(kind, 2 * 3 * m, k, seed=5 + n)
    act = np.random.default_rng(6 + n).uniform(-1, 1, size=(4, 6, n, k)).astype(np.float32)
    got = b200q.mul_mat(b200q.upload(_code(kind), (2, 3, m, k), raw), torch.from_numpy(act).cuda()).cpu().numpy()
    wv = wv.reshape(2, 3, m, k)
    for i3 in range(4):
        for i2 in range(6):
            want = act[i3, i2].astype(np.float64) @ wv[i3 // 2, i2 // 2].T
            assert _nmse(got[i3, i2], want) <= MAX_NMSE


@pytest.mark.parametrize("kind", ["f16", "bf16", "f32"])
@pytest.mark.parametrize("n", [1, 3, 12])
def test_dense_mul_mat_strided_rows(kind, n):
    """a view into a wider tensor (the KV cache): weight rows at twice their length, activation and dst columns padded"""
    import torch
    import b200q
    m, k = 50, 45
    raw, wv = _weights(kind, m, 2 * k, seed=17 + n)
    es = raw.shape[1] // (2 * k)
    act = np.random.default_rng(18 + n).uniform(-1, 1, size=(n, k + 3)).astype(np.float32)
    w_dev = torch.from_numpy(np.ascontiguousarray(raw)).cuda()
    a_dev = torch.from_numpy(act).cuda()
    out = torch.full((n, m + 5), -7.0, dtype=torch.float32, device="cuda")
    a = b200q.MulMatArgs(_code(kind), k, m, n, 1, 1, 1, 1, 2 * k * es, 2 * k * es * m, 2 * k * es * m, (k + 3) * 4, (k + 3) * 4 * n,
                         (k + 3) * 4 * n, (m + 5) * 4, (m + 5) * 4 * n, (m + 5) * 4 * n, 0)
    assert b200q.lib().b200q_mul_mat_workspace(C.byref(a)) == 0
    rc = b200q.lib().b200q_mul_mat(C.byref(a), w_dev.data_ptr(), a_dev.data_ptr(), out.data_ptr(), None, 0,
                                   C.c_void_p(torch.cuda.current_stream().cuda_stream))
    assert rc == 0, b200q.lib().b200q_last_error()
    got = out.cpu().numpy()
    want = act[:, :k].astype(np.float64) @ wv[:, :k].T
    assert _nmse(got[:, :m], want) <= MAX_NMSE
    assert (got[:, m:] == -7.0).all()  # nothing outside the rows
