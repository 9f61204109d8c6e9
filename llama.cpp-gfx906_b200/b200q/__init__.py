"""b200q — host-side mirror of the reference's operator interface for the quantized matmul hot path.

The names and argument meaning follow ggml's operators (``ggml_mul_mat`` ggml/src/ggml.c:3015-3030,
``ggml_mul_mat_id`` :3056-3080, ``ggml_get_rows``; type numbers ggml/include/ggml.h:371-411) so the parity
tests read like the reference's ``test-backend-ops`` cases.  All compute happens in ``libb200q.so``
(hand-written sm_100a CUDA behind the C-ABI of ``include/b200q.h``); PyTorch is used only to own device
memory and the CUDA stream.  There is NO CPU fallback: importing works anywhere (so the symbol/ABI tests
can run without a GPU), but every compute call raises if the library or a CUDA device is missing.
"""
from __future__ import annotations

import ctypes as C
import os
from dataclasses import dataclass
from typing import Optional

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("B200Q_LIB", os.path.join(os.path.dirname(_HERE), "lib", "libb200q.so"))

# ggml_type numbers
F32, Q4_0, Q4_1, Q5_0, Q5_1, Q8_0, Q4_K, Q5_K, Q6_K, MXFP4 = 0, 2, 3, 6, 7, 8, 12, 13, 14, 39
Q2_K, Q3_K, IQ4_NL, IQ4_XS = 10, 11, 20, 23
F16, BF16 = 1, 30  # unquantized weights (MUL_MAT / GET_ROWS): rows of plain values, any k
TYPE_NAMES = {Q4_0: "q4_0", Q4_1: "q4_1", Q5_0: "q5_0", Q5_1: "q5_1", Q8_0: "q8_0", Q4_K: "q4_K", Q5_K: "q5_K",
              Q6_K: "q6_K", MXFP4: "mxfp4", Q2_K: "q2_K", Q3_K: "q3_K", IQ4_NL: "iq4_nl", IQ4_XS: "iq4_xs",
              F16: "f16", BF16: "bf16", F32: "f32"}
BLCK = {Q4_0: (32, 18), Q4_1: (32, 20), Q5_0: (32, 22), Q5_1: (32, 24), Q8_0: (32, 34), MXFP4: (32, 17),
        Q4_K: (256, 144), Q5_K: (256, 176), Q6_K: (256, 210), F32: (1, 4), Q2_K: (256, 84), Q3_K: (256, 110),
        IQ4_NL: (32, 18), IQ4_XS: (256, 136), F16: (1, 2), BF16: (1, 2)}

FLAG_WEIGHTS_IMMUTABLE, FLAG_FORCE_GEMV, FLAG_FORCE_MMQ, FLAG_NO_PDL = 1, 2, 4, 8

EXPORTS = [
    "b200q_version", "b200q_last_error", "b200q_device_count", "b200q_device_info", "b200q_blck_size",
    "b200q_type_size", "b200q_row_size", "b200q_type_supported", "b200q_dequantize", "b200q_get_rows",
    "b200q_mul_mat_workspace", "b200q_mul_mat", "b200q_mul_mat_group", "b200q_mul_mat_group_peers", "b200q_ipc_export",
    "b200q_ipc_import", "b200q_ipc_close", "b200q_mul_mat_id_workspace", "b200q_mul_mat_id",
    "b200q_mul_mat_host", "b200q_set_device", "b200q_malloc", "b200q_free", "b200q_memset", "b200q_memcpy",
    "b200q_stream_create", "b200q_stream_destroy", "b200q_stream_synchronize", "b200q_host_alloc", "b200q_host_free",
    "b200q_list_create", "b200q_list_launch", "b200q_list_destroy", "b200q_list_info", "b200q_mul_mat_static",
    "b200q_event_create", "b200q_event_destroy", "b200q_event_record", "b200q_event_synchronize", "b200q_stream_wait_event",
    "b200q_enable_peer_access", "b200q_memcpy_peer", "b200q_max_k",
    "b200q_op_rms_norm", "b200q_op_bin", "b200q_op_scale", "b200q_op_unary", "b200q_op_glu", "b200q_op_soft_max", "b200q_op_rope",
    "b200q_graph_begin", "b200q_graph_end", "b200q_graph_launch", "b200q_graph_destroy",
    "b200q_op_cpy", "b200q_op_add_id", "b200q_op_argsort", "b200q_op_norm_fused", "b200q_op_attn_decode", "b200q_op_set_rows_pair", "b200q_op_rope_store", "b200q_op_moe_route", "b200q_op_moe_gate_weights", "b200q_scatter_rows_peers", "b200q_op_moe_weighted_sum", "b200q_mul_mat_id_bias", "b200q_op_set_rows", "b200q_mxf_probe",
]


class MulMatArgs(C.Structure):
    _fields_ = [("type", C.c_int), ("k", C.c_int64), ("m", C.c_int64), ("n", C.c_int64),
                ("ne02", C.c_int64), ("ne03", C.c_int64), ("ne12", C.c_int64), ("ne13", C.c_int64),
                ("nb01", C.c_size_t), ("nb02", C.c_size_t), ("nb03", C.c_size_t),
                ("nb11", C.c_size_t), ("nb12", C.c_size_t), ("nb13", C.c_size_t),
                ("nb1", C.c_size_t), ("nb2", C.c_size_t), ("nb3", C.c_size_t), ("flags", C.c_uint32)]


class MulMatGroupArgs(C.Structure):
    _fields_ = [("type", C.c_int), ("k", C.c_int64), ("n_mats", C.c_int), ("m", C.c_int64 * 4), ("flags", C.c_uint32)]


class MulMatIdArgs(C.Structure):
    _fields_ = [("type", C.c_int), ("k", C.c_int64), ("m", C.c_int64), ("n_expert", C.c_int64),
                ("n_used", C.c_int64), ("n_tokens", C.c_int64), ("b_ne1", C.c_int64),
                ("nb01", C.c_size_t), ("nb02", C.c_size_t), ("nb11", C.c_size_t), ("nb12", C.c_size_t),
                ("ids_nb1", C.c_size_t), ("nb1", C.c_size_t), ("nb2", C.c_size_t), ("flags", C.c_uint32)]


class ListOp(C.Structure):
    _fields_ = [("type", C.c_int), ("m", C.c_int64), ("k", C.c_int64), ("w", C.c_void_p), ("act", C.c_void_p),
                ("dst", C.c_void_p), ("flags", C.c_uint32), ("act2", C.c_void_p)]


class ListPeers(C.Structure):
    _fields_ = [("rank", C.c_int), ("world", C.c_int), ("peer_delta", C.POINTER(C.c_int64)), ("sync", C.POINTER(C.c_void_p))]


LIST_DEP = 1
LIST_SWIGLU = 2


class GetRowsArgs(C.Structure):
    _fields_ = [("type", C.c_int), ("ne00", C.c_int64), ("ne10", C.c_int64), ("ne11", C.c_int64), ("ne12", C.c_int64),
                ("nb01", C.c_size_t), ("nb02", C.c_size_t), ("nb03", C.c_size_t),
                ("nb10", C.c_size_t), ("nb11", C.c_size_t), ("nb12", C.c_size_t),
                ("nb1", C.c_size_t), ("nb2", C.c_size_t), ("nb3", C.c_size_t)]


_lib = None


def lib() -> C.CDLL:
    """Load libb200q.so (built in-tree by __graft_entry__.build()).  Fails loudly when it is missing."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(f"{LIB_PATH} not built: run `python -c 'import __graft_entry__ as g; g.build()'` "
                               "(there is no CPU fallback)")
        L = C.CDLL(LIB_PATH)
        L.b200q_last_error.restype = C.c_char_p
        L.b200q_blck_size.restype = C.c_int64
        L.b200q_max_k.restype = C.c_int64
        L.b200q_max_k.argtypes = [C.c_int]
        L.b200q_type_size.restype = C.c_size_t
        L.b200q_row_size.restype = C.c_size_t
        L.b200q_row_size.argtypes = [C.c_int, C.c_int64]
        L.b200q_mul_mat_workspace.restype = C.c_size_t
        L.b200q_mul_mat_id_workspace.restype = C.c_size_t
        L.b200q_dequantize.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p]
        L.b200q_get_rows.argtypes = [C.POINTER(GetRowsArgs), C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        L.b200q_mul_mat.argtypes = [C.POINTER(MulMatArgs), C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t,
                                    C.c_void_p]
        L.b200q_mul_mat_workspace.argtypes = [C.POINTER(MulMatArgs)]
        L.b200q_mul_mat_static.argtypes = [C.POINTER(MulMatArgs), C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        L.b200q_mul_mat_id.argtypes = [C.POINTER(MulMatIdArgs), C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                       C.c_void_p, C.c_size_t, C.c_void_p]
        L.b200q_mul_mat_id_workspace.argtypes = [C.POINTER(MulMatIdArgs)]
        L.b200q_mul_mat_host.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_int64, C.c_int64,
                                         C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p]
        L.b200q_list_create.argtypes = [C.POINTER(ListOp), C.c_int, C.POINTER(ListPeers), C.POINTER(C.c_void_p)]
        L.b200q_list_launch.argtypes = [C.c_void_p, C.c_void_p]
        L.b200q_list_destroy.argtypes = [C.c_void_p]
        L.b200q_list_info.argtypes = [C.c_void_p, C.POINTER(C.c_int)]
        _lib = L
    return _lib


def _check(rc: int, what: str):
    if rc != 0:
        raise RuntimeError(f"{what} failed ({rc}): {lib().b200q_last_error().decode()}")


def row_size(qtype: int, k: int) -> int:
    be, bb = BLCK[qtype]
    assert k % be == 0, f"k={k} not a multiple of the {TYPE_NAMES.get(qtype, qtype)} block size {be}"
    return k // be * bb


def _torch():
    import torch
    if not torch.cuda.is_available():
        raise RuntimeError("b200q needs a CUDA device (no CPU fallback)")
    return torch


def _stream():
    return C.c_void_p(_torch().cuda.current_stream().cuda_stream)


@dataclass
class QTensor:
    """A quantized ggml tensor resident in HBM: ``data`` holds the unchanged GGUF block bytes of
    ``shape = (..., m, k)`` (ggml ne = [k, m, ...]), row pitch = row_size(type, k), plus 64 bytes of slack so the
    kernels' 16-byte aligned bulk windows never leave the allocation."""
    qtype: int
    shape: tuple
    data: "object"  # torch.uint8 cuda tensor

    @property
    def k(self):
        return self.shape[-1]

    @property
    def m(self):
        return self.shape[-2]

    @property
    def nbytes(self):
        n = 1
        for s in self.shape[:-1]:
            n *= s
        return n * row_size(self.qtype, self.k)


def upload(qtype: int, shape, host_bytes: np.ndarray, device="cuda") -> QTensor:
    """ggml_backend_tensor_set of a quantized tensor: raw block bytes -> HBM."""
    torch = _torch()
    shape = tuple(int(s) for s in shape)
    n = 1
    for s in shape[:-1]:
        n *= s
    nbytes = n * row_size(qtype, shape[-1])
    hb = np.ascontiguousarray(host_bytes).view(np.uint8).reshape(-1)
    assert hb.size == nbytes, (hb.size, nbytes)
    buf = torch.zeros(nbytes + 64, dtype=torch.uint8, device=device)
    buf[:nbytes].copy_(torch.from_numpy(hb))
    return QTensor(qtype, shape, buf)


class Workspace:
    """grow-only device scratch (the plugin keeps one per backend stream)"""

    def __init__(self):
        self.buf = None

    def get(self, nbytes: int):
        torch = _torch()
        if nbytes == 0:
            return None, 0
        if self.buf is None or self.buf.numel() < nbytes:
            self.buf = torch.empty(nbytes, dtype=torch.uint8, device="cuda")
        return self.buf, self.buf.numel()


_ws = Workspace()


def dequantize(w: QTensor):
    """type_traits.to_float on the device: f32 tensor of w.shape (bit-exact vs dequantize_row_*)."""
    torch = _torch()
    n = 1
    for s in w.shape:
        n *= s
    out = torch.empty(w.shape, dtype=torch.float32, device=w.data.device)
    _check(lib().b200q_dequantize(w.qtype, w.data.data_ptr(), out.data_ptr(), n, _stream()), "b200q_dequantize")
    return out


def get_rows(w: QTensor, rows):
    """ggml_get_rows(a, b): a = w [m, k] quantized, b = i32 indices [n_idx] (or [n1, n0]) -> f32 [..., k]."""
    torch = _torch()
    assert rows.dtype == torch.int32 and rows.is_cuda
    k, m = w.k, w.m
    rs = row_size(w.qtype, k)
    r2 = rows.reshape(-1) if rows.dim() == 1 else rows
    if rows.dim() == 1:
        ne10, ne11, ne12 = rows.numel(), 1, 1
        nb10, nb11, nb12 = rows.stride(0) * 4, 0, 0
    else:
        assert rows.dim() == 2 and len(w.shape) == 3 and w.shape[0] == rows.shape[0]
        ne10, ne11, ne12 = rows.shape[1], rows.shape[0], 1
        nb10, nb11, nb12 = rows.stride(1) * 4, rows.stride(0) * 4, 0
    out = torch.empty(tuple(rows.shape) + (k,), dtype=torch.float32, device=rows.device)
    a = GetRowsArgs(w.qtype, k, ne10, ne11, ne12, rs, rs * m, 0, nb10, nb11, nb12, k * 4, k * 4 * ne10,
                    k * 4 * ne10 * ne11)
    _check(lib().b200q_get_rows(C.byref(a), w.data.data_ptr(), r2.data_ptr(), out.data_ptr(), _stream()),
           "b200q_get_rows")
    return out


def mul_mat(w: QTensor, act, out=None, flags: int = FLAG_WEIGHTS_IMMUTABLE):
    """ggml_mul_mat(a, b): a = w quantized [(ne03, ne02,) m, k]; b = act f32 [(ne13, ne12,) n, k] -> [(..,) n, m]."""
    torch = _torch()
    assert act.dtype == torch.float32 and act.is_cuda and act.stride(-1) == 1
    k, m = w.k, w.m
    assert act.shape[-1] == k
    wb = w.shape[:-2]
    ab = act.shape[:-2]
    n = act.shape[-2]
    ne02, ne03 = (wb[-1] if len(wb) >= 1 else 1), (wb[-2] if len(wb) >= 2 else 1)
    ne12, ne13 = (ab[-1] if len(ab) >= 1 else 1), (ab[-2] if len(ab) >= 2 else 1)
    if out is None:
        out = torch.empty(tuple(ab) + (n, m), dtype=torch.float32, device=act.device)
    rs = row_size(w.qtype, k)
    st = [s * 4 for s in act.stride()]
    nb11 = st[-2]
    nb12 = st[-3] if len(ab) >= 1 else nb11 * n
    nb13 = st[-4] if len(ab) >= 2 else nb12 * ne12
    a = MulMatArgs(w.qtype, k, m, n, ne02, ne03, ne12, ne13, rs, rs * m, rs * m * ne02, nb11, nb12, nb13,
                   m * 4, m * 4 * n, m * 4 * n * ne12, flags)
    need = lib().b200q_mul_mat_workspace(C.byref(a))
    ws, ws_n = _ws.get(need)
    _check(lib().b200q_mul_mat(C.byref(a), w.data.data_ptr(), act.data_ptr(), out.data_ptr(),
                               ws.data_ptr() if ws is not None else None, ws_n, _stream()), "b200q_mul_mat")
    return out


STATIC_WORKSPACE_BYTES = 16 << 20  # B200Q_STATIC_WORKSPACE_BYTES (include/b200q.h)


def mul_mat_static(w: QTensor, act, out, flags: int = FLAG_WEIGHTS_IMMUTABLE):
    """b200q_mul_mat_static: mul_mat on 2-D operands into a caller-provided ``out`` with the library's static scratch —
    neither this wrapper nor the call allocates device memory.  Shapes whose scratch exceeds STATIC_WORKSPACE_BYTES raise."""
    torch = _torch()
    assert len(w.shape) == 2 and act.dim() == 2 and act.dtype == torch.float32 and act.is_cuda and act.stride(-1) == 1
    k, m, n = w.k, w.m, act.shape[0]
    assert act.shape[1] == k and out.dtype == torch.float32 and out.is_cuda and out.shape == (n, m) and out.is_contiguous()
    rs = row_size(w.qtype, k)
    nb11 = act.stride(0) * 4
    a = MulMatArgs(w.qtype, k, m, n, 1, 1, 1, 1, rs, rs * m, rs * m, nb11, nb11 * n, nb11 * n, m * 4, m * 4 * n, m * 4 * n,
                   flags)
    _check(lib().b200q_mul_mat_static(C.byref(a), w.data.data_ptr(), act.data_ptr(), out.data_ptr(), _stream()),
           "b200q_mul_mat_static")
    return out


def mul_mat_group_ptr(ws, act, dst_ptrs, peer_deltas=(), flags: int = FLAG_WEIGHTS_IMMUTABLE):
    """mul_mat_group on raw dst device pointers (row split: slices of a b200q_malloc'ed, IPC-shared buffer); with
    peer_deltas every result row is also stored into the peers' buffers (b200q_mul_mat_group_peers)."""
    torch = _torch()
    assert 1 <= len(ws) <= 4 and len(dst_ptrs) == len(ws)
    assert act.dtype == torch.float32 and act.is_cuda and act.numel() == ws[0].k and act.is_contiguous()
    a = MulMatGroupArgs(ws[0].qtype, ws[0].k, len(ws), (C.c_int64 * 4)(*([w.m for w in ws] + [0] * (4 - len(ws)))), flags)
    wp = (C.c_void_p * len(ws))(*[w.data.data_ptr() for w in ws])
    dp = (C.c_void_p * len(ws))(*[int(d) for d in dst_ptrs])
    pd = (C.c_int64 * max(1, len(peer_deltas)))(*([int(d) for d in peer_deltas] or [0]))
    _check(lib().b200q_mul_mat_group_peers(C.byref(a), wp, C.c_void_p(act.data_ptr()), dp, len(peer_deltas), pd, _stream()),
           "b200q_mul_mat_group_peers")


def device_malloc(nbytes: int) -> int:
    p = C.c_void_p()
    _check(lib().b200q_malloc(C.byref(p), C.c_size_t(nbytes)), "b200q_malloc")
    return int(p.value)


def ipc_export(dev_ptr: int) -> bytes:
    h = C.create_string_buffer(64)
    _check(lib().b200q_ipc_export(C.c_void_p(dev_ptr), h), "b200q_ipc_export")
    return h.raw


def ipc_import(handle: bytes) -> int:
    p = C.c_void_p()
    _check(lib().b200q_ipc_import(C.c_char_p(handle), C.byref(p)), "b200q_ipc_import")
    return int(p.value)


def mul_mat_group(ws, act, outs=None, flags: int = FLAG_WEIGHTS_IMMUTABLE):
    """n = 1 ggml_mul_mat of up to 4 weight matrices of one type against the SAME activation vector (wq/wk/wv,
    ffn_gate/ffn_up) in one launch; returns the list of f32 results [1, m_g] (bit-identical to separate mul_mat calls)."""
    torch = _torch()
    assert 1 <= len(ws) <= 4 and all(len(w.shape) == 2 and w.qtype == ws[0].qtype and w.k == ws[0].k for w in ws)
    assert act.dtype == torch.float32 and act.is_cuda and act.numel() == ws[0].k and act.is_contiguous()
    if outs is None:
        outs = [torch.empty((1, w.m), dtype=torch.float32, device=act.device) for w in ws]
    a = MulMatGroupArgs(ws[0].qtype, ws[0].k, len(ws), (C.c_int64 * 4)(*([w.m for w in ws] + [0] * (4 - len(ws)))), flags)
    wp = (C.c_void_p * len(ws))(*[w.data.data_ptr() for w in ws])
    dp = (C.c_void_p * len(ws))(*[o.data_ptr() for o in outs])
    _check(lib().b200q_mul_mat_group(C.byref(a), wp, C.c_void_p(act.data_ptr()), dp, _stream()), "b200q_mul_mat_group")
    return outs


def mul_mat_id(w: QTensor, b, ids, n_used: Optional[int] = None, out=None, flags: int = FLAG_WEIGHTS_IMMUTABLE):
    """ggml_mul_mat_id(as, b, ids): as = w [n_expert, m, k]; b f32 [n_tokens, 1 | n_used, k];
    ids i32 [n_tokens, >= n_used] (may be a view of a wider tensor) -> [n_tokens, n_used, m]."""
    torch = _torch()
    assert len(w.shape) == 3 and b.dim() == 3 and ids.dim() == 2 and ids.dtype == torch.int32
    n_expert, m, k = w.shape
    n_tokens, b_ne1, _ = b.shape
    n_used = n_used or ids.shape[1]
    assert ids.stride(1) == 1 and b.stride(2) == 1
    if out is None:
        out = torch.empty((n_tokens, n_used, m), dtype=torch.float32, device=b.device)
    rs = row_size(w.qtype, k)
    a = MulMatIdArgs(w.qtype, k, m, n_expert, n_used, n_tokens, b_ne1, rs, rs * m, b.stride(1) * 4, b.stride(0) * 4,
                     ids.stride(0) * 4, m * 4, m * 4 * n_used, flags)
    need = lib().b200q_mul_mat_id_workspace(C.byref(a))
    ws, ws_n = _ws.get(need)
    _check(lib().b200q_mul_mat_id(C.byref(a), w.data.data_ptr(), b.data_ptr(), ids.data_ptr(), out.data_ptr(),
                                  ws.data_ptr() if ws is not None else None, ws_n, _stream()), "b200q_mul_mat_id")
    return out


class MatmulList:
    """A run of n = 1 ggml_mul_mat nodes executed by ONE persistent kernel launch (b200q_list_*): ``ops`` is a sequence of
    (w: QTensor [m, k], act: f32 cuda tensor [k], dst: f32 cuda tensor [m] or a raw device address, dep: bool).  ``dep``
    = the op's activations are produced by earlier ops of the list (the kernel waits for all of them, on every GPU of a
    row split).  A fifth member ``act2`` makes the op read silu(act) * act2 (B200Q_LIST_SWIGLU: ffn_down behind ffn_gate and
    ffn_up).  ``peers`` = (rank, world, peer_deltas, sync_ptrs) for row split (see include/b200q.h)."""

    def __init__(self, ops, peers=None):
        torch = _torch()
        self._keep = []
        arr = (ListOp * len(ops))()
        for i, op in enumerate(ops):
            w, act, dst, dep = op[:4]
            act2 = op[4] if len(op) > 4 else None
            assert act2 is None or (act2.dtype == torch.float32 and act2.is_cuda and act2.is_contiguous() and act2.numel() == w.k)
            assert act.dtype == torch.float32 and act.is_cuda and act.is_contiguous() and act.numel() == w.k
            dptr = dst if isinstance(dst, int) else dst.data_ptr()
            if not isinstance(dst, int):
                assert dst.dtype == torch.float32 and dst.is_cuda and dst.is_contiguous() and dst.numel() == w.m
            arr[i] = ListOp(w.qtype, w.m, w.k, w.data.data_ptr() if w.m else None, act.data_ptr(), dptr,
                            (LIST_DEP if dep else 0) | (LIST_SWIGLU if act2 is not None else 0), act2.data_ptr() if act2 is not None else None)
            self._keep.append((w, act, dst, act2))
        pp = None
        if peers is not None:
            rank, world, deltas, syncs = peers
            self._pd = (C.c_int64 * max(1, len(deltas)))(*([int(d) for d in deltas] or [0]))
            self._ps = (C.c_void_p * world)(*[int(x) for x in syncs])
            self._peers = ListPeers(rank, world, self._pd, self._ps)
            pp = C.byref(self._peers)
        h = C.c_void_p()
        _check(lib().b200q_list_create(arr, len(ops), pp, C.byref(h)), "b200q_list_create")
        self.handle = h
        info = (C.c_int * 4)()
        _check(lib().b200q_list_info(h, info), "b200q_list_info")
        self.n_launches, self.n_barrier_points, self.n_generations, self.n_slots = info[0], info[1], info[2], info[3]

    def launch(self):
        _check(lib().b200q_list_launch(self.handle, _stream()), "b200q_list_launch")

    def close(self):
        if self.handle:
            lib().b200q_list_destroy(self.handle)
            self.handle = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
