"""Synthetic workloads of BASELINE.json: the matmul lists of the models the reference's headline numbers are quoted
on (shapes from SURVEY §8a; quant mix of Q4_K_M from the reference's src/llama-quant.cpp:184-186,302-303,358-364)."""
from __future__ import annotations

from dataclasses import dataclass
from typing import List

from . import BLCK, MXFP4, Q4_K, Q6_K, Q8_0, TYPE_NAMES, row_size


@dataclass(frozen=True)
class MatMul:
    name: str
    qtype: int
    k: int
    m: int

    @property
    def weight_bytes(self) -> int:
        return self.m * row_size(self.qtype, self.k)

    def algo_bytes(self, n: int = 1) -> int:
        """algorithmic bytes of one GEMV: weights + f32 activations in + f32 result out"""
        return self.weight_bytes + 4 * n * self.k + 4 * n * self.m

    def flops(self, n: int) -> int:
        return 2 * self.m * self.k * n


def use_more_bits(i_layer: int, n_layers: int) -> bool:
    """llama-quant.cpp:185-187"""
    return i_layer < n_layers // 8 or i_layer >= 7 * n_layers // 8 or (i_layer - n_layers // 8) % 3 == 2


def llama_decode_set(n_embd: int, n_ff: int, n_kv: int, n_vocab: int, n_layer: int, mix: str = "q4_k_m",
                     layers: int | None = None, with_output: bool = True) -> List[MatMul]:
    """per-token matmul list of a Llama-architecture model (wq wk wv wo gate up down per layer + output head)"""
    ops: List[MatMul] = []
    L = n_layer if layers is None else layers
    for il in range(L):
        if mix == "q4_k_m":
            more = use_more_bits(il, n_layer)
            t_v = Q6_K if more else Q4_K
            t_down = Q6_K if more else Q4_K
            t = Q4_K
        elif mix == "q8_0":
            t = t_v = t_down = Q8_0
        elif mix == "q4_k":
            t = t_v = t_down = Q4_K
        else:
            raise ValueError(mix)
        ops += [
            MatMul(f"blk.{il}.attn_q", t, n_embd, n_embd),
            MatMul(f"blk.{il}.attn_k", t, n_embd, n_kv),
            MatMul(f"blk.{il}.attn_v", t_v, n_embd, n_kv),
            MatMul(f"blk.{il}.attn_output", t, n_embd, n_embd),
            MatMul(f"blk.{il}.ffn_gate", t, n_embd, n_ff),
            MatMul(f"blk.{il}.ffn_up", t, n_embd, n_ff),
            MatMul(f"blk.{il}.ffn_down", t_down, n_ff, n_embd),
        ]
    if with_output:
        ops.append(MatMul("output", Q6_K if mix in ("q4_k_m", "q4_k") else Q8_0, n_embd, n_vocab))
    return ops


def llama3_8b(mix: str = "q4_k_m", layers: int | None = None, with_output: bool = True) -> List[MatMul]:
    return llama_decode_set(4096, 14336, 1024, 128256, 32, mix, layers, with_output)


def llama3_70b(mix: str = "q4_k", layers: int | None = None, with_output: bool = True) -> List[MatMul]:
    return llama_decode_set(8192, 28672, 1024, 128256, 80, mix, layers, with_output)


def describe(ops: List[MatMul]) -> dict:
    by = {}
    for o in ops:
        key = f"{TYPE_NAMES[o.qtype]} {o.k}->{o.m}"
        by[key] = by.get(key, 0) + 1
    return {"n_matmuls": len(ops), "weight_bytes": sum(o.weight_bytes for o in ops), "shapes": by}


def synth_weights_device(op: MatMul, seed: int, rows: int | None = None, row0: int = 0):
    """random-init quantized weights of the op's architecture generated directly in HBM: random payload bits with finite
    fp16 / e8m0 scales (no dataset or checkpoint exists offline).  Returns a b200q.QTensor of rows [row0, row0+rows)."""
    import torch
    from . import QTensor
    m = op.m if rows is None else rows
    be, bb = BLCK[op.qtype]
    nblk = m * (op.k // be)
    g = torch.Generator(device="cuda")
    g.manual_seed(seed * 1000003 + row0)
    raw = torch.randint(0, 256, (nblk, bb), dtype=torch.uint8, device="cuda", generator=g)

    def half(lo, hi, n=nblk):
        v = torch.rand(n, device="cuda", generator=g) * (hi - lo) + lo
        return v.to(torch.float16).view(torch.uint8).reshape(n, 2)

    t = op.qtype
    if t == Q4_K or TYPE_NAMES[t] == "q5_K":
        raw[:, 0:2] = half(0.0, 0.02)
        raw[:, 2:4] = half(0.0, 0.02)
    elif t == Q6_K:
        raw[:, 208:210] = half(-0.01, 0.01)
    elif t == MXFP4:
        raw[:, 0] = torch.randint(115, 128, (nblk,), dtype=torch.uint8, device="cuda", generator=g)
    elif TYPE_NAMES[t] in ("q4_1", "q5_1"):
        raw[:, 0:2] = half(0.0, 0.5)
        raw[:, 2:4] = half(-1.0, 1.0)
    else:
        raw[:, 0:2] = half(-0.5, 0.5)
    buf = torch.zeros(nblk * bb + 64, dtype=torch.uint8, device="cuda")
    buf[:nblk * bb] = raw.reshape(-1)
    del raw
    return QTensor(t, (m, op.k), buf)


def row_partition(m: int, world: int, align: int = 128) -> list:
    """row split of an [m, k] weight matrix over `world` ranks: equal cumulative fractions rounded down to `align`
    rows, the last rank takes the remainder (the reference's get_row_split, ggml-cuda.cu:760-786, with equal
    tensor_split on identical GPUs).  Returns [(row0, nrows)] per rank."""
    bounds = [((m * r) // world) // align * align for r in range(world)] + [m]
    return [(bounds[r], bounds[r + 1] - bounds[r]) for r in range(world)]
