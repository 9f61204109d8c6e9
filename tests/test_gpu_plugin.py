"""The drop-in boundary under test: libggml-b200.so (the ggml backend plugin) driven by the reference's OWN code.

  * test_backend_ops_*: the unmodified reference harness (oracle/_ref/**/test-backend-ops, built from
    /root/reference/tests/test-backend-ops.cpp) runs its MUL_MAT / MUL_MAT_ID / GET_ROWS cases against the plugin
    (GGML_BACKEND_PATH, ggml-backend-reg.cpp:588-592) with ggml-cpu as its oracle: NMSE <= 5e-4 for the matmuls
    (test-backend-ops.cpp:3106-3108, 3209-3211), 1e-7 for GET_ROWS (:948-950).  Asserted: rc 0, zero FAIL lines, and at
    least the number of supported + passing cases the previous round logged.
  * test_inprocess_*: the plugin dlopen'ed INSIDE this process through ggml_backend_load (tests/plugin_driver.cpp),
    used directly (alloc / set / graph_compute / get, workspace growth inside one graph) and through
    ggml_backend_sched next to the CPU backend (supports_op, buffer types, cross-backend copies).
"""
import ctypes as C
import os
import re
import subprocess

import numpy as np
import pytest

import refs

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PLUGIN = os.path.join(ROOT, "llama.cpp-gfx906_b200", "lib", "libggml-b200.so")

pytestmark = pytest.mark.gpu


def _need_ref():
    d = refs.ref_dir()
    if d is None or not os.path.exists(PLUGIN):
        pytest.skip("oracle/_ref (reference build) or the plugin is not present")
    return d


def _run_tbo(op, timeout=900):
    d = _need_ref()
    exe = os.path.join(d, "test-backend-ops")
    env = dict(os.environ, GGML_BACKEND_PATH=PLUGIN)
    p = subprocess.run([exe, "test", "-o", op], env=env, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True,
                       timeout=timeout)
    out = re.sub(r"\x1b\[[0-9;]*m", "", p.stdout)
    lines = [l for l in out.splitlines() if l.lstrip().startswith(op + "(")]
    ok = sum(1 for l in lines if l.rstrip().endswith("OK"))
    fail = [l for l in lines if "FAIL" in l]
    unsup = sum(1 for l in lines if "not supported" in l)
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    with open(os.path.join(ROOT, "gpurun_out", f"tbo_{op}.txt"), "w") as f:
        f.write(out)
    assert "Backend B2000" in out, out[-2000:]
    return p.returncode, ok, fail, unsup, out


# supported-and-passing counts logged in profiles/r02_test_backend_ops_*_v2.txt (round 1: 43 / 243 / 194); a regression in supports_op or in a
# kernel shows up as a lower count or a FAIL line
# the operators around the matmuls (csrc/ops.cu): NMSE <= 1e-7 against ggml-cpu, the harness' default (test-backend-ops.cpp:948-950)
LAYER_OPS = [("RMS_NORM", 8), ("ADD", 33), ("SUB", 26), ("MUL", 26), ("DIV", 26), ("SCALE", 2), ("SILU", 2), ("GELU", 2),
             ("GELU_QUICK", 2), ("GELU_ERF", 2), ("RELU", 2), ("SIGMOID", 2), ("TANH", 2), ("NEG", 2), ("ABS", 2), ("EXP", 2),
             ("REGLU", 12), ("GEGLU", 12), ("SWIGLU", 12), ("SWIGLU_OAI", 8), ("GEGLU_ERF", 12), ("GEGLU_QUICK", 12),
             ("SOFT_MAX", 199), ("ROPE", 176), ("CPY", 55), ("CONT", 10), ("DUP", 6), ("SET_ROWS", 61), ("ADD_ID", 34), ("ARGSORT", 6)]


@pytest.mark.parametrize("op,min_ok", [("GET_ROWS", 65), ("MUL_MAT", 870), ("MUL_MAT_ID", 270)] + LAYER_OPS)
def test_backend_ops_against_plugin(op, min_ok):
    rc, ok, fail, unsup, out = _run_tbo(op)
    assert not fail, fail[:5]
    assert rc == 0, out[-2000:]
    assert ok >= min_ok, (ok, unsup)
    assert re.search(r"Backend B2000: OK", out)


_pd = None


def _driver():
    global _pd
    if _pd is None:
        d = _need_ref()
        # a host that uses ggml links ggml-base into its global scope; the plugin resolves ggml_backend_buffer_init & co. from
        # there (it is not linked against any ggml library itself), so do the same in this Python process
        for lib in ("libggml-base.so", "libggml-cpu.so", "libggml.so"):
            C.CDLL(os.path.join(d, lib), mode=C.RTLD_GLOBAL)
        L = C.CDLL(os.path.join(d, "plugin_driver.so"), mode=C.RTLD_GLOBAL)
        L.pd_load.argtypes = [C.c_char_p]
        L.pd_device_name.restype = C.c_char_p
        L.pd_proc_address.restype = C.c_void_p
        L.pd_proc_address.argtypes = [C.c_char_p]
        L.pd_mul_mat_seq.argtypes = [C.c_int, C.c_int, C.c_int64, C.c_int64, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p,
                                     C.c_void_p]
        L.pd_sched_mlp.argtypes = [C.c_int, C.c_int, C.c_int64, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64,
                                   C.c_void_p, C.c_void_p, C.c_void_p]
        L.pd_mul_mat_split.argtypes = [C.c_int, C.c_void_p, C.c_int, C.c_int64, C.c_int64, C.c_void_p, C.c_int64, C.c_void_p,
                                       C.c_void_p, C.c_void_p]
        L.pd_async_roundtrip.argtypes = [C.c_int, C.c_int, C.c_void_p, C.c_int64, C.c_void_p]
        L.pd_big_upload.argtypes = [C.c_int, C.c_void_p, C.c_int64, C.c_void_p]
        L.pd_decode_run.argtypes = [C.c_int, C.c_int, C.c_int, C.c_int64] + [C.c_void_p] * 6
        L.pd_layer_replay.argtypes = [C.c_int, C.c_int, C.c_int64, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p]
        L.pd_last_us.restype = C.c_double
        L.pd_llama_layer.argtypes = [C.c_int, C.c_int, C.c_int, C.c_int, C.POINTER(C.c_double), C.POINTER(C.c_int)]
        L.pd_llama_layer_bias.argtypes = L.pd_llama_layer.argtypes
        L.pd_llama_layer_vtrans.argtypes = [C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.POINTER(C.c_double), C.POINTER(C.c_int)]
        L.pd_moe_block.argtypes = [C.c_int, C.c_int, C.POINTER(C.c_double), C.POINTER(C.c_int)]
        n = L.pd_load(PLUGIN.encode())
        assert n >= 1, f"ggml_backend_load({PLUGIN}) registered {n} devices"
        _pd = L
    return _pd


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


def test_inprocess_load_and_registry():
    L = _driver()
    assert L.pd_device_name(0) == b"B2000"
    assert L.pd_proc_address(b"ggml_backend_get_features")
    assert L.pd_proc_address(b"ggml_backend_split_buffer_type")  # llama's -sm row hook (src/llama-model.cpp:372-373)
    with open("/proc/self/maps") as f:
        assert "libggml-b200.so" in f.read()


@pytest.mark.parametrize("t", [refs.Q4_K, refs.Q8_0], ids=lambda t: refs.NAMES[t])
def test_inprocess_mul_mat_workspace_growth(t):
    """one graph_compute, four MUL_MAT nodes with 1, 40, 300 and 700 columns: GEMV, then the tensor-core kernel whose
    workspace has to grow twice while earlier nodes are still in flight on the stream"""
    L = _driver()
    m, k = 384, 1024
    ns = np.array([1, 40, 300, 700], dtype=np.int64)
    w = refs.make_weights(t, m, k, seed=31 + t)
    act = refs.uniform_weights(int(ns.max()), k, seed=32)
    out = np.zeros(int(ns.sum()) * m, dtype=np.float32)
    rc = L.pd_mul_mat_seq(0, t, m, k, _p(w), _p(ns), len(ns), _p(act), _p(out))
    assert rc == 0, rc
    off = 0
    for n in ns:
        n = int(n)
        want = refs.ref_mul_mat(t, w, act[:n])[0]
        got = out[off:off + n * m].reshape(n, m)
        assert refs.nmse(got, want) <= 5e-4, n
        off += n * m


def test_inprocess_scheduler_with_cpu_backend():
    """W2 . silu(W1 . x) through ggml_backend_sched over [B200, CPU] equals the CPU backend alone (NMSE <= 5e-4)"""
    L = _driver()
    t, k, m, n = refs.Q4_K, 1024, 2048, 3
    w1 = refs.make_weights(t, m, k, seed=41)
    w2 = refs.make_weights(t, k, m, seed=42)
    x = refs.uniform_weights(n, k, seed=43)
    y = np.zeros((n, k), dtype=np.float32)
    y_cpu = np.zeros((n, k), dtype=np.float32)
    splits = C.c_int(0)
    rc = L.pd_sched_mlp(0, t, k, m, _p(w1), _p(w2), _p(x), n, _p(y), _p(y_cpu), C.byref(splits))
    assert rc == 0, rc
    assert np.isfinite(y).all() and np.abs(y_cpu).max() > 0
    assert refs.nmse(y, y_cpu) <= 5e-4
    assert splits.value >= 1


def _n_devices():
    return _driver().pd_load(PLUGIN.encode())


@pytest.mark.parametrize("t,m,k,n", [(refs.Q4_K, 4096, 4096, 1), (refs.Q6_K, 1024, 4096, 1), (refs.Q4_K, 1000, 1024, 5),
                                     (refs.Q8_0, 2048, 2048, 64), (refs.Q4_K, 300, 512, 1)],
                         ids=lambda v: refs.NAMES.get(v, str(v)) if isinstance(v, int) and v in refs.NAMES and v >= 0 else str(v))
def test_inprocess_split_buffer_equals_plain(t, m, k, n):
    """MUL_MAT on a weight in ggml_backend_split_buffer_type equals the same MUL_MAT on the device's plain buffer BIT FOR BIT
    (the kernels' summation order does not depend on how the rows are distributed).  With one device the split degenerates
    to a single slice; with >= 2 the slices live on different GPUs and the peers store their rows into the main device's
    dst over NVLink.  300 rows: a device can end up without rows (rounding to 128)."""
    L = _driver()
    nd = _n_devices()
    w = refs.make_weights(t, m, k, seed=51 + t + m)
    act = refs.uniform_weights(n, k, seed=52)
    for split in ([None] if nd == 1 else [None, [3.0, 1.0] + [0.0] * 14]):
        a = np.zeros((n, m), dtype=np.float32)
        b = np.zeros((n, m), dtype=np.float32)
        ts = None if split is None else (C.c_float * 16)(*split)
        rc = L.pd_mul_mat_split(0, ts, t, m, k, _p(w), n, _p(act), _p(a), _p(b))
        assert rc == 0, rc
        assert np.abs(b).max() > 0
        assert a.tobytes() == b.tobytes(), (split, float(np.abs(a - b).max()))
        want = refs.ref_mul_mat(t, w, act)[0]
        assert refs.nmse(a, want) <= 5e-4


def test_inprocess_split_buffer_other_main_device():
    """main device 1: its backend runs the op, device 0 is the peer (needs 2 GPUs)"""
    L = _driver()
    if _n_devices() < 2:
        pytest.skip("needs 2 GPUs")
    t, m, k, n = refs.Q4_K, 2048, 4096, 1
    w = refs.make_weights(t, m, k, seed=61)
    act = refs.uniform_weights(n, k, seed=62)
    a = np.zeros((n, m), dtype=np.float32)
    b = np.zeros((n, m), dtype=np.float32)
    assert L.pd_mul_mat_split(1, None, t, m, k, _p(w), n, _p(act), _p(a), _p(b)) == 0
    assert a.tobytes() == b.tobytes()


def test_inprocess_events_async_copies_host_buffer():
    """pinned host buffer type -> set_tensor_async -> cpy_tensor_async to a second backend (a second GPU when there is one)
    -> get_tensor_async + event_synchronize"""
    L = _driver()
    n = 1 << 20
    x = refs.uniform_weights(1, n, seed=71).reshape(-1)
    for dev_b in sorted({0, _n_devices() - 1}):
        y = np.zeros(n, dtype=np.float32)
        rc = L.pd_async_roundtrip(0, dev_b, _p(x), n, _p(y))
        assert rc == 0, (rc, dev_b)
        assert y.tobytes() == x.tobytes()


def test_inprocess_staged_upload():
    """set_tensor of 40 MB + 5 bytes from pageable memory: three pinned staging pieces, the last one ragged"""
    L = _driver()
    nbytes = 40 * (1 << 20) + 5
    src = np.random.default_rng(81).integers(0, 256, nbytes, dtype=np.uint8)
    back = np.zeros(nbytes, dtype=np.uint8)
    assert L.pd_big_upload(0, _p(src), nbytes, _p(back)) == 0


@pytest.mark.parametrize("t,tv", [(refs.Q4_K, refs.Q6_K), (refs.Q8_0, refs.Q8_0), (refs.Q4_K, refs.Q5_K)],
                         ids=["q4_K+q6_K", "q8_0", "q4_K+q5_K(not in the list kernel)"])
def test_inprocess_decode_run_as_list(t, tv):
    """q / k / v on one vector and o = W4 q in one graph: the plugin launches them as ONE persistent list (a barrier point in
    front of o); the results equal the node-by-node path (GGML_B200_NO_LIST) bit for bit and the reference within NMSE 5e-4.
    A run with a type the list kernel does not serve (q5_K) falls back to single launches."""
    L = _driver()
    k = 2048  # q6_K rows of 1680 bytes: a multiple of 16, so the per-node path takes the same ring kernel family as the list
    ws = [refs.make_weights(tt, k, k, seed=91 + i) for i, tt in enumerate([t, t, tv, t])]
    x = refs.uniform_weights(1, k, seed=95)
    got = np.zeros((4, k), dtype=np.float32)
    one = np.zeros((4, k), dtype=np.float32)
    os.environ.pop("GGML_B200_NO_LIST", None)
    assert L.pd_decode_run(0, t, tv, k, _p(ws[0]), _p(ws[1]), _p(ws[2]), _p(ws[3]), _p(x), _p(got)) == 0
    os.environ["GGML_B200_NO_LIST"] = "1"
    try:
        assert L.pd_decode_run(0, t, tv, k, _p(ws[0]), _p(ws[1]), _p(ws[2]), _p(ws[3]), _p(x), _p(one)) == 0
    finally:
        os.environ.pop("GGML_B200_NO_LIST", None)
    refs_ = [refs.ref_mul_mat(tt, ws[i], x)[0] for i, tt in enumerate([t, t, tv])] + [refs.ref_mul_mat(t, ws[3], one[0:1])[0]]
    detail = [(float(np.abs(got[i] - one[i]).max()), refs.nmse(got[i], refs_[i]), refs.nmse(one[i], refs_[i])) for i in range(4)]
    assert got.tobytes() == one.tobytes(), detail
    for i, tt in enumerate([t, t, tv]):
        assert refs.nmse(got[i], refs.ref_mul_mat(tt, ws[i], x)[0]) <= 5e-4
    assert refs.nmse(got[3], refs.ref_mul_mat(t, ws[3], got[0:1])[0]) <= 5e-4


@pytest.mark.parametrize("t", [refs.Q4_K, refs.Q8_0], ids=lambda t: refs.NAMES[t])
def test_inprocess_graph_replay(t):
    """a 20-node layer-like graph computed six times on changing inputs: pass 1 runs eagerly, pass 2 is recorded as a CUDA
    graph, passes 3-6 are replays (ggml-b200.cpp backend_graph_compute; the reference's cudaGraph path, ggml-cuda.cu:2876-3100).
    Every pass must equal the all-eager run (GGML_B200_NO_GRAPHS) bit for bit, and the passes must differ from each other."""
    L = _driver()
    k, layers, reps = 1024, 4, 6
    wa = refs.make_weights(t, k, k, seed=131)
    wb = refs.make_weights(t, k, k, seed=132)
    x = refs.uniform_weights(1, k, seed=133)
    got = np.zeros((reps, k), dtype=np.float32)
    eager = np.zeros((reps, k), dtype=np.float32)
    os.environ.pop("GGML_B200_NO_GRAPHS", None)
    stats = C.CFUNCTYPE(None, C.POINTER(C.c_int64), C.POINTER(C.c_int64))(L.pd_proc_address(b"ggml_backend_b200_graph_stats"))
    rec0, rep0, rec1, rep1 = C.c_int64(), C.c_int64(), C.c_int64(), C.c_int64()
    stats(C.byref(rec0), C.byref(rep0))
    assert L.pd_layer_replay(0, t, k, layers, _p(wa), _p(wb), _p(x), reps, _p(got)) == 0
    stats(C.byref(rec1), C.byref(rep1))
    assert rec1.value - rec0.value == 1 and rep1.value - rep0.value == reps - 2, (rec1.value - rec0.value, rep1.value - rep0.value)
    os.environ["GGML_B200_NO_GRAPHS"] = "1"
    try:
        assert L.pd_layer_replay(0, t, k, layers, _p(wa), _p(wb), _p(x), reps, _p(eager)) == 0
    finally:
        os.environ.pop("GGML_B200_NO_GRAPHS", None)
    assert np.isfinite(got).all()
    assert got.tobytes() == eager.tobytes(), [float(np.abs(got[r] - eager[r]).max()) for r in range(reps)]
    for r in range(1, reps):
        assert not np.array_equal(got[r], got[r - 1])


def test_inprocess_graph_replay_is_not_slower():
    """a 32-layer decode-like graph (160 nodes): time per graph_compute + synchronize, replayed vs eager; printed for the log"""
    L = _driver()
    t, k, layers, reps = refs.Q4_K, 4096, 32, 40
    wa = refs.make_weights(t, k, k, seed=141)
    wb = refs.make_weights(t, k, k, seed=142)
    x = refs.uniform_weights(1, k, seed=143)
    out = np.zeros((1, k), dtype=np.float32)
    os.environ.pop("GGML_B200_NO_GRAPHS", None)
    assert L.pd_layer_replay(0, t, k, layers, _p(wa), _p(wb), _p(x), reps, _p(out)) == 0
    us_graph = L.pd_last_us()
    os.environ["GGML_B200_NO_GRAPHS"] = "1"
    try:
        assert L.pd_layer_replay(0, t, k, layers, _p(wa), _p(wb), _p(x), reps, _p(out)) == 0
    finally:
        os.environ.pop("GGML_B200_NO_GRAPHS", None)
    us_eager = L.pd_last_us()
    print(f"\n[graph replay] 32-layer decode-like graph (160 nodes, q4_K 4096x4096): eager {us_eager:.1f} us, replay {us_graph:.1f} us")
    assert us_graph <= us_eager * 1.10


@pytest.mark.parametrize("t", [refs.Q4_K, refs.Q8_0, refs.Q6_K], ids=lambda t: refs.NAMES[t])
@pytest.mark.parametrize("n_tokens,n_past", [(1, 37), (1, 0), (5, 60), (17, 5), (40, 0)])
def test_inprocess_llama_layer_stays_on_device(t, n_tokens, n_past):
    """one whole Llama decoder layer as llama.cpp builds it without flash attention — norm, q/k/v, RoPE, KV
    store (SET_ROWS into an f16 cache), K^T q on the cache view, masked soft_max, V kq, wo, residual, norm, SwiGLU FFN, residual —
    every node is accepted by the plugin's supports_op (nothing bounces to the CPU) and the layer output equals the reference
    CPU backend's on the same weights, inputs and pre-filled cache (decode, small batch and prompt shapes)."""
    L = _driver()
    nmse, unsup = C.c_double(-1.0), C.c_int(-1)
    before = _fusion_stats(L)
    rc = L.pd_llama_layer(0, t, n_tokens, n_past, C.byref(nmse), C.byref(unsup))
    assert rc == 0 and unsup.value == 0, (rc, unsup.value)
    fired = {k: v - before[k] for k, v in _fusion_stats(L).items()}
    print(f"\n[llama layer] {refs.NAMES[t]} n_tokens={n_tokens} n_past={n_past}: nmse vs CPU backend {nmse.value:.2e}; fusions {fired}")
    assert 0.0 <= nmse.value <= 1e-3
    if n_tokens == 1 and not any(os.environ.get(k) for k in ("GGML_B200_NO_FUSE", "GGML_B200_NO_LIST", "GGML_B200_NO_HOIST", "GGML_B200_NO_SCRATCH")):
        # the decode step is what the fusions are for: they must have fired on the graph the CPU backend was compared with
        # (two norms, the attention core, q / k / v and gate / up as lists, rope + both cache stores as one launch)
        assert fired["norm"] >= 2 and fired["attn"] >= 1 and fired["list"] >= 2 and fired["rope_store"] >= 1, fired


@pytest.mark.parametrize("n_tokens,n_past", [(1, 37), (1, 0), (5, 60), (40, 0)])
@pytest.mark.parametrize("t", [refs.Q8_0, refs.Q4_K], ids=lambda t: refs.NAMES[t])
def test_inprocess_layer_with_attention_biases(t, n_tokens, n_past):
    """the same layer with q / k / v / o biases (gpt-oss, Qwen2: an ADD behind each matmul).  At decode the ADDs in front of the
    ROPEs and of the V store are part of the one launch behind the q / k / v list (b200q_op_rope_store with biases) and read
    hoisted results from the scratch; the output equals the reference CPU backend's."""
    L = _driver()
    nmse, unsup = C.c_double(-1.0), C.c_int(-1)
    before = _fusion_stats(L)
    rc = L.pd_llama_layer_bias(0, t, n_tokens, n_past, C.byref(nmse), C.byref(unsup))
    assert rc == 0 and unsup.value == 0, (rc, unsup.value)
    fired = {k: v - before[k] for k, v in _fusion_stats(L).items()}
    print(f"\n[llama layer + biases] {refs.NAMES[t]} n_tokens={n_tokens} n_past={n_past}: nmse vs CPU backend {nmse.value:.2e}; fusions {fired}")
    assert 0.0 <= nmse.value <= 1e-3
    if n_tokens == 1 and not any(os.environ.get(k) for k in ("GGML_B200_NO_FUSE", "GGML_B200_NO_LIST", "GGML_B200_NO_HOIST", "GGML_B200_NO_SCRATCH")):
        assert fired["rope_store"] >= 1 and fired["list"] >= 2, fired


@pytest.mark.parametrize("bias", [0, 1])
@pytest.mark.parametrize("n_tokens,n_past", [(1, 37), (1, 0), (5, 60), (40, 0)])
def test_inprocess_layer_with_transposed_v_cache(n_tokens, n_past, bias):
    """the layer as llama-bench -fa 0 really runs it: the V cache is kept transposed (llama-kv-cache.cpp v_trans: a SET_ROWS of
    one-element rows with an index per element writes it, the attention reads it through a strided view, no CONT).  This is the
    layout the fused rope + cache stores launch and the fused attention core meet in the benchmark; checked against the CPU
    backend with and without attention biases."""
    L = _driver()
    nmse, unsup = C.c_double(-1.0), C.c_int(-1)
    before = _fusion_stats(L)
    rc = L.pd_llama_layer_vtrans(0, refs.Q4_K, n_tokens, n_past, bias, C.byref(nmse), C.byref(unsup))
    assert rc == 0 and unsup.value == 0, (rc, unsup.value)
    fired = {k: v - before[k] for k, v in _fusion_stats(L).items()}
    print(f"\n[llama layer, transposed V] n_tokens={n_tokens} n_past={n_past} bias={bias}: nmse vs CPU backend {nmse.value:.2e}; fusions {fired}")
    assert 0.0 <= nmse.value <= 1e-3
    if n_tokens == 1 and not any(os.environ.get(k) for k in ("GGML_B200_NO_FUSE", "GGML_B200_NO_LIST", "GGML_B200_NO_HOIST", "GGML_B200_NO_SCRATCH")):
        assert fired["rope_store"] >= 1 and fired["attn"] >= 1, fired


FUSIONS = ("norm", "attn", "rope_store", "set_rows_pair", "moe_sum", "id_bias", "list", "scratch_read", "scratch_copy", "glu_list", "route")


def _fusion_stats(L):
    """the plugin's counters of how often each graph_dispatch fusion fired (proc address ggml_backend_b200_fusion_stats)"""
    L.pd_proc_address.restype = C.c_void_p
    fn = C.CFUNCTYPE(C.c_int, C.POINTER(C.c_int64), C.c_int)(L.pd_proc_address(b"ggml_backend_b200_fusion_stats"))
    buf = (C.c_int64 * len(FUSIONS))()
    assert fn(buf, len(FUSIONS)) == len(FUSIONS)
    return dict(zip(FUSIONS, buf))


@pytest.mark.parametrize("n_tokens", [1, 7, 40, 200])
def test_inprocess_gpt_oss_moe_block_stays_on_device(n_tokens):
    """the MoE feed-forward block of gpt-oss (src/llama-graph.cpp build_moe_ffn, SOFTMAX_WEIGHT gating) with MXFP4 experts: f32
    router matmul + bias, top-k (a view of ARGSORT), GET_ROWS, soft_max over the selection, MUL_MAT_ID x3 with ADD_ID biases,
    SWIGLU_OAI, weighting and the sum over the used experts — every node accepted by supports_op, output equal to the
    reference CPU backend's (the routing is discrete: equal selections are part of what the NMSE checks)."""
    L = _driver()
    nmse, unsup = C.c_double(-1.0), C.c_int(-1)
    before = _fusion_stats(L)
    rc = L.pd_moe_block(0, n_tokens, C.byref(nmse), C.byref(unsup))
    assert rc == 0 and unsup.value == 0, (rc, unsup.value)
    fired = {k: v - before[k] for k, v in _fusion_stats(L).items()}
    print(f"\n[gpt-oss moe block] n_tokens={n_tokens}: nmse vs CPU backend {nmse.value:.2e}; fusions {fired}")
    assert 0.0 <= nmse.value <= 1e-3
    if not os.environ.get("GGML_B200_NO_FUSE"):  # router (bias, top-k, soft_max over the selection) and combine as one launch each
        assert fired["route"] >= 2 and fired["moe_sum"] >= 1, fired  # both halves: ADD + ARGSORT, GET_ROWS + SOFT_MAX


def test_inprocess_split_buffer_only_serves_mul_mat():
    """llama.cpp chooses each weight's buffer type by asking supports_op for the operator it is used in; with -sm row a norm
    weight (MUL / ADD) or an embedding (GET_ROWS) offered in the split buffer type must be declined, MUL_MAT accepted
    (ggml-cuda.cu:3265-3283).  Found by llama-bench -sm row: the norm weights landed in the split buffer and the fused norm read
    a slice table as an address."""
    L = _driver()
    assert L.pd_split_supports(0) == 0
