"""Test-side access to the oracle (oracle/liboracle.so, our C restatement) and, when present, to the unmodified
reference compiled by oracle/Makefile (oracle/_ref/ref_driver.so).  TEST INFRASTRUCTURE ONLY."""
import ctypes as C
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")

F32, Q4_0, Q4_1, Q5_0, Q5_1, Q8_0, Q8_1, Q4_K, Q5_K, Q6_K, Q8_K, MXFP4 = 0, 2, 3, 6, 7, 8, 9, 12, 13, 14, 15, 39
Q2_K, Q3_K, IQ4_NL, IQ4_XS = 10, 11, 20, 23
WEIGHT_TYPES = [Q4_0, Q4_1, Q5_0, Q5_1, Q8_0, MXFP4, Q4_K, Q5_K, Q6_K, Q2_K, Q3_K, IQ4_NL, IQ4_XS]
HEADLINE_TYPES = [Q4_0, Q8_0, MXFP4, Q4_K, Q6_K]
NAMES = {Q4_0: "q4_0", Q4_1: "q4_1", Q5_0: "q5_0", Q5_1: "q5_1", Q8_0: "q8_0", Q8_1: "q8_1", Q4_K: "q4_K",
         Q5_K: "q5_K", Q6_K: "q6_K", Q8_K: "q8_K", MXFP4: "mxfp4", Q2_K: "q2_K", Q3_K: "q3_K", IQ4_NL: "iq4_nl",
         IQ4_XS: "iq4_xs"}
BLCK = {Q4_0: (32, 18), Q4_1: (32, 20), Q5_0: (32, 22), Q5_1: (32, 24), Q8_0: (32, 34), Q8_1: (32, 36),
        MXFP4: (32, 17), Q4_K: (256, 144), Q5_K: (256, 176), Q6_K: (256, 210), Q8_K: (256, 292), Q2_K: (256, 84),
        Q3_K: (256, 110), IQ4_NL: (32, 18), IQ4_XS: (256, 136)}

_oracle = None
_ref = None


def row_size(t, k):
    be, bb = BLCK[t]
    assert k % be == 0
    return k // be * bb


def oracle():
    global _oracle
    if _oracle is None:
        so = os.path.join(ORACLE_DIR, "liboracle.so")
        src = os.path.join(ORACLE_DIR, "oracle.c")
        if not os.path.exists(so) or os.path.getmtime(so) < os.path.getmtime(src):
            subprocess.check_call(["make", "-C", ORACLE_DIR, "oracle"], stdout=subprocess.DEVNULL)
        L = C.CDLL(so)
        L.orc_fp32_to_fp16.restype = C.c_uint16
        L.orc_fp32_to_fp16.argtypes = [C.c_float]
        L.orc_fp16_to_fp32.restype = C.c_float
        L.orc_fp16_to_fp32.argtypes = [C.c_uint16]
        L.orc_dequantize_row.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_int64]
        L.orc_quantize_row.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_int64]
        L.orc_vec_dot.argtypes = [C.c_int, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p]
        L.orc_mul_mat.argtypes = [C.c_int, C.c_void_p, C.c_size_t, C.c_void_p, C.c_void_p, C.c_int64, C.c_int64, C.c_int64]
        L.orc_mul_mat_id.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p] + [C.c_int64] * 7
        L.orc_get_rows.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_int64]
        _oracle = L
    return _oracle


def _cpu_flags():
    try:
        with open("/proc/cpuinfo") as f:
            for line in f:
                if line.startswith("flags"):
                    return set(line.split(":", 1)[1].split())
    except OSError:
        pass
    return set()


def ref_dir():
    """directory of the reference build matching this host's CPU, or None"""
    base = os.path.join(ORACLE_DIR, "_ref")
    flags = _cpu_flags()
    if {"avx512f", "avx512bw", "avx512vl", "avx512dq", "avx512cd", "avx512vbmi", "avx512_vnni"} <= flags and \
            os.path.exists(os.path.join(base, "avx512", "ref_driver.so")):
        return os.path.join(base, "avx512")
    if {"avx2", "fma", "f16c", "bmi2"} <= flags and os.path.exists(os.path.join(base, "ref_driver.so")):
        return base
    return None


def ref():
    """the reference's own code behind oracle/ref_driver.cpp, or None when oracle/_ref was not built"""
    global _ref
    if _ref is None:
        d = ref_dir()
        if d is None:
            return None
        L = C.CDLL(os.path.join(d, "ref_driver.so"))
        L.ref_cpu_features.restype = C.c_char_p
        L.ref_row_size.restype = C.c_size_t
        L.ref_row_size.argtypes = [C.c_int, C.c_int64]
        L.ref_quantize.restype = C.c_size_t
        L.ref_quantize.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_int64, C.c_int64]
        L.ref_dequantize.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_int64]
        L.ref_act_quantize.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_int64]
        L.ref_vec_dot.argtypes = [C.c_int, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p]
        L.ref_mul_mat.restype = C.c_double
        L.ref_mul_mat.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_int64, C.c_int64, C.c_int, C.c_int]
        L.ref_mul_mat_batched.restype = C.c_double
        L.ref_mul_mat_batched.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_void_p] + [C.c_int64] * 7 + [C.c_int, C.c_int]
        L.ref_mul_mat_id.restype = C.c_double
        L.ref_mul_mat_id.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p] + [C.c_int64] * 7 + [C.c_int, C.c_int]
        L.ref_get_rows.restype = C.c_double
        L.ref_get_rows.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_int64, C.c_int64, C.c_int]
        _ref = L
    return _ref


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


# ---------------------------------------------------------------------------------------------------------
# data
# ---------------------------------------------------------------------------------------------------------
def random_blocks(t, nrows, k, seed):
    """random but numerically sane quantized rows: every payload bit pattern occurs, scales are finite fp16"""
    rng = np.random.default_rng(seed)
    be, bb = BLCK[t]
    nb = nrows * (k // be)
    raw = rng.integers(0, 256, size=(nb, bb), dtype=np.uint8)

    def half(lo, hi):
        return rng.uniform(lo, hi, size=nb).astype(np.float16).view(np.uint8).reshape(nb, 2)

    if t in (Q4_0, Q5_0, Q8_0):
        raw[:, 0:2] = half(-0.5, 0.5)
    elif t == IQ4_NL:
        raw[:, 0:2] = half(-0.01, 0.01)
    elif t == IQ4_XS:
        raw[:, 0:2] = half(-0.001, 0.001)
    elif t == Q2_K:
        raw[:, 80:82] = half(0.0, 0.05)
        raw[:, 82:84] = half(0.0, 0.05)
    elif t == Q3_K:
        raw[:, 108:110] = half(-0.01, 0.01)
    elif t in (Q4_1, Q5_1):
        raw[:, 0:2] = half(0.0, 0.5)
        raw[:, 2:4] = half(-1.0, 1.0)
    elif t == MXFP4:
        raw[:, 0] = rng.integers(115, 128, size=nb, dtype=np.uint8)
    elif t in (Q4_K, Q5_K):
        raw[:, 0:2] = half(0.0, 0.02)
        raw[:, 2:4] = half(0.0, 0.02)
    elif t == Q6_K:
        raw[:, 208:210] = half(-0.01, 0.01)
    return raw.reshape(nrows, -1)


def uniform_weights(nrows, k, seed):
    """the reference harness' distribution: U(-1, 1) (tests/test-backend-ops.cpp:42-131), fixed seed"""
    return np.random.default_rng(seed).uniform(-1.0, 1.0, size=(nrows, k)).astype(np.float32)


def quantize_weights(t, w):
    """quantize f32 rows with the reference (ggml_quantize_chunk) when available, else with the oracle
    restatement (q4_0/q8_0/mxfp4 only); returns uint8 [nrows, row_size] or None"""
    nrows, k = w.shape
    out = np.zeros((nrows, row_size(t, k)), dtype=np.uint8)
    w = np.ascontiguousarray(w, dtype=np.float32)
    R = ref()
    if R is not None:
        R.ref_quantize(t, _p(w), _p(out), nrows, k)
        return out
    if t in (Q4_0, Q8_0, MXFP4):
        O = oracle()
        for i in range(nrows):
            assert O.orc_quantize_row(t, _p(w[i]), _p(out[i]), k) == 0
        return out
    return None


def make_weights(t, nrows, k, seed):
    """quantized U(-1,1) weights when a quantizer is available for `t`, else random valid blocks"""
    q = quantize_weights(t, uniform_weights(nrows, k, seed))
    return q if q is not None else random_blocks(t, nrows, k, seed)


# ---------------------------------------------------------------------------------------------------------
# oracle ops (numpy in / numpy out)
# ---------------------------------------------------------------------------------------------------------
def orc_dequantize(t, rows, k):
    rows = np.ascontiguousarray(rows)
    out = np.empty((rows.shape[0], k), dtype=np.float32)
    O = oracle()
    for i in range(rows.shape[0]):
        assert O.orc_dequantize_row(t, _p(rows[i]), _p(out[i]), k) == 0
    return out


def orc_mul_mat(t, w, act):
    """w uint8 [m, row_size], act f32 [n, k] -> f32 [n, m]"""
    w = np.ascontiguousarray(w)
    act = np.ascontiguousarray(act, dtype=np.float32)
    m, n, k = w.shape[0], act.shape[0], act.shape[1]
    out = np.empty((n, m), dtype=np.float32)
    assert oracle().orc_mul_mat(t, _p(w), w.shape[1], _p(act), _p(out), m, k, n) == 0
    return out


def orc_mul_mat_id(t, w, b, ids, n_used):
    """w uint8 [n_expert, m, row_size]; b f32 [n_tokens, b_ne1, k]; ids i32 [n_tokens, stride] -> [n_tokens, n_used, m]"""
    w = np.ascontiguousarray(w)
    b = np.ascontiguousarray(b, dtype=np.float32)
    ids = np.ascontiguousarray(ids, dtype=np.int32)
    n_expert, m = w.shape[0], w.shape[1]
    n_tokens, b_ne1, k = b.shape
    out = np.empty((n_tokens, n_used, m), dtype=np.float32)
    rc = oracle().orc_mul_mat_id(t, _p(w), _p(b), _p(ids), _p(out), m, k, n_expert, n_used, n_tokens, b_ne1, ids.shape[1])
    assert rc == 0, rc
    return out


def ref_mul_mat(t, w, act, threads=0, repeat=1):
    R = ref()
    w = np.ascontiguousarray(w)
    act = np.ascontiguousarray(act, dtype=np.float32)
    m, n, k = w.shape[0], act.shape[0], act.shape[1]
    out = np.empty((n, m), dtype=np.float32)
    secs = R.ref_mul_mat(t, _p(w), _p(act), _p(out), m, k, n, threads, repeat)
    assert secs >= 0
    return out, secs


def nmse(a, b):
    """normalised mean squared error as test-backend-ops defines it (tests/test-backend-ops.cpp:133-177 callers)"""
    a = np.asarray(a, dtype=np.float64).ravel()
    b = np.asarray(b, dtype=np.float64).ravel()
    den = float(np.sum(b * b))
    return float(np.sum((a - b) ** 2)) / den if den > 0 else float(np.sum((a - b) ** 2))
