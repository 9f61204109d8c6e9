#!/usr/bin/env python3
"""SURVEY H3 / VERDICT item 7(i), the precision half: what would MXFP4 weights x activations give on the block-scaled
tensor-core path (tcgen05.mma kind::mxf8f6f4: e2m1 A with UE8M0 block-32 scales, e4m3 B with UE8M0 block-32 scales)?
Pure numerics, emulated exactly on the CPU (every product of an e2m1 code point, an e4m3 value and two powers of two is
exact in fp64; the MMA accumulates in fp32, which adds ~1e-7 relative): the NMSE against the fp64 product of the bit-exact
dequantized weights with the f32 activations, the quantity test-backend-ops gates at 5e-4.
Variants:  (a) activations as ONE e4m3 term per element;  (b) hi + lo two-term split (two MMAs per tile);
(c) what mmq.cu does today: d*q rounded to bf16 x activations rounded to bf16;  (d) int8 q8_0 activations (ggml's own)."""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import refs  # noqa: E402


def e4m3(x):
    return torch.from_numpy(np.ascontiguousarray(x, dtype=np.float32)).to(torch.float8_e4m3fn).to(torch.float32).numpy().astype(np.float64)


def ue8m0_scale(amax, target):
    """smallest power of two s with amax / s <= target"""
    e = np.ceil(np.log2(np.maximum(amax, 1e-30) / target))
    return np.exp2(e)


def block_e4m3(x):
    """per-32 UE8M0 scale + e4m3 codes; returns the dequantized values"""
    n, k = x.shape
    b = x.reshape(n, k // 32, 32).astype(np.float64)
    s = ue8m0_scale(np.abs(b).max(axis=2, keepdims=True), 448.0)
    return (e4m3(b / s) * s).reshape(n, k)


def main():
    rows = []
    for k, m, n, dist in [(2880, 256, 64, "uniform"), (2880, 256, 64, "normal"), (4096, 256, 64, "uniform")]:
        w = refs.uniform_weights(m, k, 11)
        rng = np.random.default_rng(12)
        act = rng.uniform(-1, 1, (n, k)).astype(np.float32) if dist == "uniform" else rng.standard_normal((n, k)).astype(np.float32)
        wq = refs.quantize_weights(refs.MXFP4, w)
        wd = refs.orc_dequantize(refs.MXFP4, wq, k).astype(np.float64)  # exact fp4 code points x 2^e
        exact = act.astype(np.float64) @ wd.T
        a1 = block_e4m3(act)
        lo = block_e4m3(act.astype(np.float64) - a1)
        bf = lambda x: torch.from_numpy(np.ascontiguousarray(x, dtype=np.float32)).to(torch.bfloat16).to(torch.float32).numpy().astype(np.float64)
        q8 = np.empty_like(act, dtype=np.float64)
        ab = act.reshape(n, k // 32, 32).astype(np.float64)
        d = np.abs(ab).max(axis=2, keepdims=True) / 127.0
        q8 = (np.round(ab / np.where(d > 0, d, 1)) * d).reshape(n, k)
        res = {
            "(a) e4m3 x1": refs.nmse(a1 @ wd.T, exact),
            "(b) e4m3 hi+lo": refs.nmse((a1 + lo) @ wd.T, exact),
            "(c) bf16 x bf16 (mmq.cu)": refs.nmse(bf(act) @ bf(wd).T, exact),
            "(d) int8 per-32 (q8_0)": refs.nmse(q8 @ wd.T, exact),
        }
        rows.append((k, m, n, dist, res))
    for k, m, n, dist, res in rows:
        print(f"k={k} m={m} n={n} activations {dist}:")
        for name, v in res.items():
            print(f"    {name:28s} NMSE {v:.3e}   {'PASS' if v <= 2.5e-4 else ('pass (no margin)' if v <= 5e-4 else 'FAIL')} (gate 5e-4, margin 2.5e-4)")


if __name__ == "__main__":
    main()
