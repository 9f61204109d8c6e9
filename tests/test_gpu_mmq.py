"""GPU parity of the tensor-core (tcgen05) prompt-processing path, n > 8, through the C-ABI.
   * against the oracle (ggml-cpu arithmetic restated, oracle/oracle.c) at sizes it finishes in seconds: NMSE <= 5e-4,
     the reference's MUL_MAT criterion (test-backend-ops.cpp:3106-3108)
   * at BASELINE sizes (Llama-3-8B shapes, n = 512) against an fp64 product of the bit-exact dequantized weights
     (b200q.dequantize is checked bit-for-bit against dequantize_row_* in test_gpu_parity.py) — the MMQ path
     rounds both operands to bf16 once, so its NMSE against that product must be ~1e-5, far inside the gate.
"""
import numpy as np
import pytest

import refs

pytestmark = pytest.mark.gpu
MAX_NMSE = 5e-4


def _run(t, m, k, n, seed, flags=0):
    import b200q
    import torch
    w = refs.make_weights(t, m, k, seed=seed)
    act = refs.uniform_weights(n, k, seed=seed + 1)
    wt = b200q.upload(t, (m, k), w)
    got = b200q.mul_mat(wt, torch.from_numpy(act).cuda(), flags=flags).cpu().numpy()
    return w, act, wt, got


@pytest.mark.parametrize("t", refs.WEIGHT_TYPES, ids=lambda t: refs.NAMES[t])
@pytest.mark.parametrize("shape", [(16, 256, 9), (16, 256, 16), (129, 1024, 33), (300, 512, 257), (128, 2048, 512),
                                   (257, 256, 40)])
def test_mmq_vs_oracle(t, shape):
    m, k, n = shape
    import b200q
    # n <= 16 would take two passes of the 8-column GEMV (api.cu use_mmq): force the tensor-core kernel
    w, act, _, got = _run(t, m, k, n, seed=700 + t + m, flags=b200q.FLAG_FORCE_MMQ)
    want = refs.orc_mul_mat(t, w, act)
    assert got.shape == want.shape
    assert np.isfinite(got).all()
    assert refs.nmse(got, want) <= MAX_NMSE


@pytest.mark.parametrize("t", [refs.MXFP4, refs.Q8_0, refs.Q4_0], ids=lambda t: refs.NAMES[t])
def test_mmq_ragged_k(t):
    """k = 2880 (gpt-oss): 45 k-chunks of 64, last W stage ragged (64 of 256 elements); 1530-byte mxfp4 rows"""
    m, k, n = 200, 2880, 48
    w, act, _, got = _run(t, m, k, n, seed=800 + t)
    want = refs.orc_mul_mat(t, w, act)
    assert refs.nmse(got, want) <= MAX_NMSE


@pytest.mark.parametrize("t", [refs.Q4_K, refs.Q8_0], ids=lambda t: refs.NAMES[t])
def test_mmq_forced_small_n_matches_gemv(t):
    """the same n <= 8 problem through both kernels (FORCE_MMQ vs GEMV) agrees at the quantization-noise level"""
    import b200q
    m, k, n = 384, 1024, 4
    _, _, _, a = _run(t, m, k, n, seed=900 + t, flags=b200q.FLAG_FORCE_MMQ)
    _, _, _, b = _run(t, m, k, n, seed=900 + t, flags=b200q.FLAG_FORCE_GEMV)
    assert refs.nmse(a, b) <= MAX_NMSE


@pytest.mark.parametrize("t", [refs.Q4_0, refs.Q4_K, refs.Q8_0], ids=lambda t: refs.NAMES[t])
@pytest.mark.parametrize("bs,nr", [((3, 1), (1, 1)), ((2, 2), (2, 2))])
def test_mmq_batched_broadcast(t, bs, nr):
    import b200q
    import torch
    m, k, n = 16, 256, 12
    ne02, ne03 = bs
    ne12, ne13 = ne02 * nr[0], ne03 * nr[1]
    w = refs.make_weights(t, ne03 * ne02 * m, k, seed=500 + t)
    act = refs.uniform_weights(ne13 * ne12 * n, k, seed=501).reshape(ne13, ne12, n, k)
    want = np.empty((ne13, ne12, n, m), dtype=np.float32)
    w4 = w.reshape(ne03, ne02, m, -1)
    for i13 in range(ne13):
        for i12 in range(ne12):
            want[i13, i12] = refs.orc_mul_mat(t, w4[i13 // nr[1], i12 // nr[0]], act[i13, i12])
    got = b200q.mul_mat(b200q.upload(t, (ne03, ne02, m, k), w), torch.from_numpy(act).cuda()).cpu().numpy()
    assert refs.nmse(got, want) <= MAX_NMSE


@pytest.mark.parametrize("t,m,k", [(refs.Q8_0, 4096, 4096), (refs.Q8_0, 14336, 4096), (refs.Q8_0, 4096, 14336),
                                   (refs.Q4_K, 14336, 4096), (refs.Q6_K, 4096, 14336), (refs.Q4_0, 4096, 4096),
                                   (refs.MXFP4, 2880, 2880)],
                         ids=lambda v: refs.NAMES.get(v, str(v)) if isinstance(v, int) and v in refs.NAMES else str(v))
def test_mmq_baseline_sizes(t, m, k):
    """pp512 at Llama-3-8B / gpt-oss shapes against the fp64 product of the bit-exact dequantized weights"""
    import b200q
    import torch
    n = 512
    w = np.tile(refs.make_weights(t, 512, k, seed=1000 + t), ((m + 511) // 512, 1))[:m]
    act = refs.uniform_weights(n, k, seed=1001)
    wt = b200q.upload(t, (m, k), w)
    a = torch.from_numpy(act).cuda()
    got = b200q.mul_mat(wt, a)
    ref = (a.double() @ b200q.dequantize(wt).double().T)
    err = float(((got.double() - ref) ** 2).sum() / (ref ** 2).sum())
    assert torch.isfinite(got).all()
    assert err <= 5e-5, err
    # the tiled rows repeat every 512: identical weight rows must give identical bits (fixed accumulation order)
    if m >= 1024:
        assert torch.equal(got[:, :512], got[:, 512:1024])


@pytest.mark.parametrize("t", [refs.Q4_0, refs.Q8_0, refs.MXFP4, refs.Q4_K, refs.Q6_K, refs.Q5_1], ids=lambda t: refs.NAMES[t])
@pytest.mark.parametrize("n_tokens,n_expert,n_used,bcast,m,k", [(40, 8, 4, True, 257, 512), (300, 6, 2, False, 128, 256),
                                                                (9, 4, 4, True, 64, 1024), (130, 32, 4, True, 96, 2880)])
def test_moe_through_mmq(t, n_tokens, n_expert, n_used, bcast, m, k):
    """MUL_MAT_ID with many (token, slot) rows: rows grouped per expert on the device, each expert's rows through the
    tensor-core kernel; against the oracle's mul_mat_id (ids a strided view, experts with zero / >256 rows)"""
    import b200q
    import torch
    if k % refs.BLCK[t][0]:
        pytest.skip("k not a multiple of the block size")
    w = refs.make_weights(t, n_expert * m, k, seed=900 + t)
    b_ne1 = 1 if bcast else n_used
    b = refs.uniform_weights(n_tokens * b_ne1, k, seed=901).reshape(n_tokens, b_ne1, k)
    rng = np.random.default_rng(902 + n_tokens)
    # skewed routing: expert 0 is popular (more than one 256-column tile at 300 tokens), the last expert never chosen
    ids_full = np.stack([rng.permutation(n_expert - 1) for _ in range(n_tokens)]).astype(np.int32)
    for r in range(0, n_tokens, 2):
        j = int(np.where(ids_full[r] == 0)[0][0])
        ids_full[r, j], ids_full[r, 0] = ids_full[r, 0], 0
    ids_full = np.concatenate([ids_full, np.zeros((n_tokens, 1), np.int32)], axis=1)  # a wider tensor: stride n_expert
    want = refs.orc_mul_mat_id(t, w.reshape(n_expert, m, -1), b, ids_full, n_used)
    ids_dev = torch.from_numpy(ids_full).cuda()[:, :n_used]
    wt = b200q.upload(t, (n_expert, m, k), w)
    got = b200q.mul_mat_id(wt, torch.from_numpy(b).cuda(), ids_dev, n_used=n_used)
    got2 = b200q.mul_mat_id(wt, torch.from_numpy(b).cuda(), ids_dev, n_used=n_used)
    assert torch.equal(got, got2)  # reproducible
    got = got.cpu().numpy()
    assert np.isfinite(got).all()
    assert refs.nmse(got, want) <= MAX_NMSE
