"""Row-split mode: every weight matrix is sharded by output rows over the GPUs of one box, each rank (one process per
GPU) multiplies its own rows, and the result slices are all-gathered over NVLink with NCCL.

Mirrors the reference's row split (``-sm row``): ``get_row_split`` ggml/src/ggml-cuda/ggml-cuda.cu:760-786 (cumulative
``tensor_split`` fractions rounded down to a row granule), the split buffer's ``set_tensor`` that uploads each device's
rows only (:883-920), and ``ggml_cuda_op_mul_mat`` (:1436-1754) which broadcasts the activations, runs the per-device
kernel and copies the dst slices back to the main device.  Differences, by design:
  * activations are kept REPLICATED on every rank (all-gather semantics), so the reference's per-op broadcast of the
    quantized activations disappears;
  * the slices travel with one NCCL all-gather enqueued on the compute stream (graph-capturable), not with per-device
    cudaMemcpyPeerAsync + events through the main device.
Per-row arithmetic is identical to the single-GPU kernels (fixed summation order), so a row-split result equals the
single-GPU result bit for bit — that is what tests/test_gpu_rowsplit.py checks.

The collective and the compute kernel are injectable (``compute=``, ``group=``) so that the host logic (partition,
ragged gather, layout) is covered by world_size-2 gloo tests on CPU; the defaults are the CUDA kernels + NCCL.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Callable, List, Optional, Sequence, Tuple

import numpy as np

from . import QTensor, row_size, upload

ROW_GRANULE = 128  # the MMQ tile height (the reference rounds to its mmq_y, ggml-cuda.cu:771-779)


def row_split(m: int, world: int, tensor_split: Optional[Sequence[float]] = None,
              granule: int = ROW_GRANULE) -> List[Tuple[int, int]]:
    """[(row0, nrows)] per rank.  Rank r owns rows [round_down(m*s_r), round_down(m*s_{r+1})) with s the cumulative
    normalised ``tensor_split`` (equal fractions by default); the last rank always ends at m (get_row_split)."""
    if world < 1:
        raise ValueError("world must be >= 1")
    if tensor_split is None:
        tensor_split = [1.0] * world
    if len(tensor_split) != world or any(f < 0 for f in tensor_split) or sum(tensor_split) <= 0:
        raise ValueError("tensor_split must hold one non-negative fraction per rank")
    total = float(sum(tensor_split))
    cum, acc = [], 0.0
    for f in tensor_split:
        cum.append(acc / total)
        acc += f
    lows = [int(m * c) // granule * granule for c in cum]
    lows[0] = 0
    bounds = lows + [m]
    for r in range(world):  # monotone even for degenerate fractions
        bounds[r + 1] = max(bounds[r + 1], bounds[r])
    return [(bounds[r], bounds[r + 1] - bounds[r]) for r in range(world)]


@dataclass
class RowSplitWeight:
    """this rank's shard of an [m, k] quantized matrix"""
    qtype: int
    m: int
    k: int
    rank: int
    world: int
    parts: List[Tuple[int, int]]       # (row0, nrows) of every rank
    shard: Optional[QTensor]           # rows [row0, row0 + nrows) of this rank, resident in HBM (None if nrows == 0)

    @property
    def row0(self):
        return self.parts[self.rank][0]

    @property
    def nrows(self):
        return self.parts[self.rank][1]


def shard_bytes(qtype: int, m: int, k: int, host_bytes: np.ndarray, row0: int, nrows: int) -> np.ndarray:
    """the raw GGUF block bytes of rows [row0, row0+nrows) (what the split buffer's set_tensor copies to one device)"""
    rs = row_size(qtype, k)
    hb = np.ascontiguousarray(host_bytes).view(np.uint8).reshape(m, rs)
    return hb[row0:row0 + nrows]


def upload_row_split(qtype: int, shape, host_bytes: np.ndarray, rank: int, world: int,
                     tensor_split: Optional[Sequence[float]] = None, uploader: Callable = upload) -> RowSplitWeight:
    """split-buffer set_tensor: only this rank's rows go to this rank's HBM (contiguous 2-D tensors only, as the
    reference requires, ggml-cuda.cu:1485-1488)."""
    if len(shape) != 2:
        raise ValueError("row split needs a contiguous 2-D weight tensor")
    m, k = int(shape[0]), int(shape[1])
    parts = row_split(m, world, tensor_split)
    row0, nrows = parts[rank]
    shard = uploader(qtype, (nrows, k), shard_bytes(qtype, m, k, host_bytes, row0, nrows)) if nrows > 0 else None
    return RowSplitWeight(qtype, m, k, rank, world, parts, shard)


def _default_compute(shard: QTensor, act, out):
    from . import mul_mat
    return mul_mat(shard, act, out=out)


def mul_mat_row_split(w: RowSplitWeight, act, out=None, group=None, compute: Callable = _default_compute,
                      dist=None):
    """dst[n, m] = act[n, k] x W[m, k]^T with W sharded by rows; every rank ends up with the full dst.
    ``act`` is replicated (same values on every rank).  The all-gather runs on the current stream."""
    import torch
    if dist is None:
        import torch.distributed as dist
    n = act.shape[-2] if act.dim() >= 2 else 1
    act2 = act.reshape(n, w.k)
    if out is None:
        out = torch.empty((n, w.m), dtype=torch.float32, device=act.device)
    equal = len({p[1] for p in w.parts}) == 1
    if w.world == 1:
        compute(w.shard, act2, out)
        return out
    if equal and n == 1:
        # the common decode case: the slices are contiguous pieces of dst, one plain all-gather, no staging
        part = out.view(-1)[w.row0:w.row0 + w.nrows]
        compute(w.shard, act2, part.view(1, w.nrows))
        dist.all_gather_into_tensor(out.view(-1), part, group=group)
        return out
    # general case (n > 1 or ragged split): dst slices are strided 2-D blocks -> gather dense [n, max_rows] blocks,
    # then place them ("transposed staging buffer")
    max_rows = max(p[1] for p in w.parts)
    mine = torch.zeros((n, max_rows), dtype=torch.float32, device=act.device)
    if w.nrows > 0:
        tmp = torch.empty((n, w.nrows), dtype=torch.float32, device=act.device)
        compute(w.shard, act2, tmp)
        mine[:, :w.nrows] = tmp
    stage = torch.empty((w.world * n, max_rows), dtype=torch.float32, device=act.device)
    dist.all_gather_into_tensor(stage, mine, group=group)
    stage = stage.view(w.world, n, max_rows)
    for r, (r0, nr) in enumerate(w.parts):
        if nr > 0:
            out[:, r0:r0 + nr] = stage[r, :, :nr]
    return out


class PeerDst:
    """Row split without a collective per matmul: every rank owns one b200q_malloc'ed result buffer of the same layout,
    maps the buffers of its peers (cudaIpc), and the decode kernel stores each result row into all of them in its
    epilogue (b200q_mul_mat_group_peers; NVLink peer stores).  ``fence()`` is the stream-ordered cross-GPU completion
    point (one tiny NCCL all-reduce): after it every rank's buffer holds every rank's rows.  Replaces the reference's
    per-op copy of the dst slices to the main device (ggml-cuda.cu:1700-1740)."""

    def __init__(self, n_floats: int, rank: int, world: int, dist=None, group=None):
        import torch
        from . import device_malloc, ipc_export, ipc_import
        if dist is None:
            import torch.distributed as dist
        self.rank, self.world, self.n, self.dist, self.group = rank, world, int(n_floats), dist, group
        self.ptr = device_malloc(self.n * 4 + 256)
        handles = [None] * world
        dist.all_gather_object(handles, ipc_export(self.ptr), group=group)
        self.peer_ptrs = [self.ptr if r == rank else ipc_import(handles[r]) for r in range(world)]
        self.deltas = [self.peer_ptrs[r] - self.ptr for r in range(world) if r != rank]
        self._flag = torch.zeros(1, device="cuda")
        # a torch view of the local buffer (no ownership: the memory lives until the process exits)
        self.tensor = _as_tensor(self.ptr, self.n)

    def addr(self, float_offset: int) -> int:
        return self.ptr + 4 * int(float_offset)

    def fence(self):
        self.dist.all_reduce(self._flag, group=self.group)


class PeerSync:
    """the sync blocks of a row-split persistent list (include/b200q.h, b200q_list_peers): every rank owns one
    zero-initialised block and maps the blocks of its peers; ``ptrs[r]`` is this process's pointer to rank r's block.
    The kernel publishes "this rank is complete up to barrier point b" into its slot of every block (NVLink, release at
    system scope) — per-matmul-group arrival flags instead of a collective."""

    BYTES = 2 << 20  # one whole allocation granule, so the IPC mapping covers exactly this buffer

    def __init__(self, rank: int, world: int, dist=None, group=None):
        import ctypes as C
        import torch
        from . import device_malloc, ipc_export, ipc_import, lib
        if dist is None:
            import torch.distributed as dist
        self.ptr = device_malloc(self.BYTES)
        assert lib().b200q_memset(C.c_void_p(self.ptr), 0, C.c_size_t(self.BYTES), None) == 0
        torch.cuda.synchronize()
        handles = [None] * world
        dist.all_gather_object(handles, ipc_export(self.ptr), group=group)
        self.ptrs = [self.ptr if r == rank else ipc_import(handles[r]) for r in range(world)]


def _as_tensor(ptr: int, n_floats: int):
    import torch

    class _Holder:
        pass

    h = _Holder()
    h.__cuda_array_interface__ = {"shape": (int(n_floats),), "typestr": "<f4", "data": (int(ptr), False), "version": 3,
                                  "strides": None}
    return torch.as_tensor(h, device="cuda")


def mul_mat_row_split_peer(ws: Sequence[RowSplitWeight], act, dst: PeerDst, float_offsets: Sequence[int]):
    """dst[off_g + row] for the rows of this rank's shards of up to 4 matrices that share ``act`` (n = 1), stored into
    every rank's buffer by the kernel itself.  Call ``dst.fence()`` before another GPU's results are read."""
    from . import mul_mat_group_ptr
    live = [(w, off) for w, off in zip(ws, float_offsets) if w.nrows > 0]
    if not live:
        return
    mul_mat_group_ptr([w.shard for w, _ in live], act.reshape(-1), [dst.addr(off + w.row0) for w, off in live], dst.deltas)
