"""Expert-parallel MUL_MAT_ID across the GPUs of one box.  NOT in the reference: ggml-cuda asserts that expert
tensors are never split (ggml-cuda.cu:2094), so this is a new design, kept next to the row split it resembles.

    * the experts of ``as`` [n_expert, m, k] are sharded by id: rank r holds experts [lo_r, hi_r) — the point is memory:
      a gpt-oss-120B-class expert tensor does not fit one GPU;
    * activations ``b`` and the routing ``ids`` are replicated (they are on every rank anyway: the router runs replicated), so
      the "dispatch" all-to-all of the usual expert parallelism costs nothing;
    * every rank computes the (token, slot) pairs whose expert it owns with the ordinary kernels (b200q_mul_mat_id on the
      compacted pairs, local expert ids) and stores the rows it computed into EVERY rank's result buffer over NVLink
      (b200q_scatter_rows_peers) — the "combine" is peer stores, no all-to-all and no NCCL on the data path;
    * one stream-ordered cross-GPU fence (PeerDst.fence) completes the step.

The host logic (partition, compaction, index arithmetic) is plain and is tested on CPU with gloo (tests/test_rowsplit_cpu.py);
the device path on 2 GPUs (tests/test_gpu_rowsplit.py)."""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass

import numpy as np


def expert_range(n_expert: int, rank: int, world: int):
    """contiguous shards, the first ``n_expert % world`` ranks take one more"""
    q, r = divmod(int(n_expert), int(world))
    lo = rank * q + min(rank, r)
    return lo, lo + q + (1 if rank < r else 0)


def owned_pairs(ids: np.ndarray, lo: int, hi: int):
    """ids [n_tokens, n_used] -> (pair index t * n_used + s of every pair routed to an expert in [lo, hi), its local expert id),
    in ascending pair order (the order fixes the rows of the compacted problem, so results do not depend on timing)"""
    flat = np.asarray(ids).reshape(-1)
    sel = np.nonzero((flat >= lo) & (flat < hi))[0].astype(np.int32)
    return sel, (flat[sel] - lo).astype(np.int32)


@dataclass
class ExpertShard:
    """this rank's experts of one tensor: ``w`` = QTensor [hi - lo, m, k] (only these are resident)"""
    w: "object"
    lo: int
    hi: int
    n_expert: int
    rank: int
    world: int


def shard_experts(qtype: int, host_rows: np.ndarray, n_expert: int, m: int, k: int, rank: int, world: int) -> ExpertShard:
    """upload the experts this rank owns; host_rows = raw block bytes of the whole tensor [n_expert * m, row bytes]"""
    from . import upload
    lo, hi = expert_range(n_expert, rank, world)
    rows = np.ascontiguousarray(host_rows.reshape(n_expert, m, -1)[lo:hi]).reshape((hi - lo) * m, -1)
    return ExpertShard(upload(qtype, (max(hi - lo, 1), m, k), rows if hi > lo else np.zeros((m, host_rows.shape[-1]), np.uint8)), lo, hi, n_expert, rank,
                       world)


def mul_mat_id_ep(shard: ExpertShard, b, ids, dst, fence=True):
    """ggml_mul_mat_id over sharded experts.  b f32 cuda [n_tokens, 1 | n_used, k] and ids i32 cuda [n_tokens, n_used], the
    same on every rank; ``dst`` = PeerDst of n_tokens * n_used * m floats: after the call (and its fence) every rank's
    dst.tensor holds the full result [n_tokens, n_used, m].  Rows owned by nobody (ids outside 0..n_expert) stay untouched."""
    import torch
    from . import _check, _stream, lib, mul_mat_id
    n_tokens, b_ne1, k = b.shape
    n_used = ids.shape[1]
    m = shard.w.shape[1]
    # the compaction needs the pair count on the host to size the launch: one small D2H of the routing (n_tokens * n_used ints)
    sel, local = owned_pairs(ids.cpu().numpy(), shard.lo, shard.hi)
    if sel.size:
        sel_d = torch.from_numpy(sel).to(b.device)
        tok = torch.div(sel_d, n_used, rounding_mode="floor")
        src_rows = tok if b_ne1 == 1 else sel_d                      # row of b (flattened [n_tokens * b_ne1, k]) each pair reads
        bp = b.reshape(n_tokens * b_ne1, k).index_select(0, src_rows.long()).reshape(sel.size, 1, k).contiguous()
        out = mul_mat_id(shard.w, bp, torch.from_numpy(local).to(b.device).reshape(sel.size, 1))     # [P, 1, m]
        deltas = (C.c_int64 * max(1, len(dst.deltas)))(*([int(d) for d in dst.deltas] or [0]))
        _check(lib().b200q_scatter_rows_peers(C.c_void_p(out.data_ptr()), C.c_void_p(sel_d.data_ptr()), C.c_void_p(dst.ptr), C.c_int64(sel.size),
                                              C.c_int64(m), C.c_int64(m), len(dst.deltas), deltas, _stream()), "b200q_scatter_rows_peers")
    if fence:
        dst.fence()
    return dst.tensor[: n_tokens * n_used * m].view(n_tokens, n_used, m)
