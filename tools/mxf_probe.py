#!/usr/bin/env python3
"""drives b200q_mxf_probe: builds the shared-memory images of a 128 x N x 128 block-scaled problem (e2m1 A, e4m3 B, UE8M0
block-32 scales), runs tcgen05.mma kind::mxf8f6f4.block_scale on one CTA and compares with an exact CPU emulation; then times
the instruction.  usage: tools/mxf_probe.py [N=128|256]"""
import ctypes as C
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "llama.cpp-gfx906_b200"), os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)
import torch  # noqa: E402

import b200q  # noqa: E402

E2M1 = np.array([0, 0.5, 1, 1.5, 2, 3, 4, 6, -0.0, -0.5, -1, -1.5, -2, -3, -4, -6], dtype=np.float64)


def sw128_image(rows_bytes):
    """rows_bytes: [R][128] uint8 (8 chunks of 16 bytes per row) -> K-major 128-byte-swizzle tile image"""
    R = rows_bytes.shape[0]
    img = np.zeros((R, 8, 16), dtype=np.uint8)
    src = rows_bytes.reshape(R, 8, 16)
    for r in range(R):
        for c in range(8):
            img[r, c ^ (r & 7)] = src[r, c]
    return img.reshape(-1)


def sf_image(sf):
    """sf: [R][4] uint8, R a multiple of 128 -> blocks of 32 x 16 bytes: byte [l][4 c + j] = sf[128 b + 32 c + l][j]"""
    out = []
    for b in range(sf.shape[0] // 128):
        blk = np.zeros((32, 16), dtype=np.uint8)
        for c in range(4):
            for l in range(32):
                blk[l, 4 * c:4 * c + 4] = sf[128 * b + 32 * c + l]
        out.append(blk.reshape(-1))
    return np.concatenate(out)


def main():
    N = int(sys.argv[1]) if len(sys.argv) > 1 else 128
    rng = np.random.default_rng(5)
    a_code = rng.integers(0, 16, (128, 128), dtype=np.uint8)
    sfa = rng.integers(120, 131, (128, 4), dtype=np.uint8)
    b_f = rng.uniform(-4, 4, (N, 128)).astype(np.float32)
    b_e4 = torch.from_numpy(b_f).to(torch.float8_e4m3fn)
    b_bytes = b_e4.view(torch.uint8).numpy()
    b_val = b_e4.to(torch.float32).numpy().astype(np.float64)
    sfb = rng.integers(122, 129, (N, 4), dtype=np.uint8)
    # reference
    a_val = E2M1[a_code]
    want = np.zeros((128, N))
    for j in range(4):
        blk = a_val[:, 32 * j:32 * j + 32] @ b_val[:, 32 * j:32 * j + 32].T
        want += blk * np.exp2(sfa[:, j].astype(np.float64) - 127)[:, None] * np.exp2(sfb[:, j].astype(np.float64) - 127)[None, :]
    L = b200q.lib()
    L.b200q_mxf_probe.argtypes = [C.c_void_p] * 5 + [C.c_int, C.c_int, C.c_void_p, C.c_void_p]
    results = {}
    for nib_lo_first in (True, False):
        for data_first in (True, False):
            rows = np.zeros((128, 8, 16), dtype=np.uint8)
            for c in range(8):
                e = a_code[:, 16 * c:16 * c + 16]
                packed = (e[:, 0::2] | (e[:, 1::2] << 4)) if nib_lo_first else (e[:, 1::2] | (e[:, 0::2] << 4))
                if data_first:
                    rows[:, c, :8] = packed
                else:
                    rows[:, c, 8:] = packed
            a_img = torch.from_numpy(sw128_image(rows.reshape(128, 128))).cuda()
            b_img = torch.from_numpy(sw128_image(b_bytes)).cuda()
            sfa_img = torch.from_numpy(sf_image(sfa)).cuda()
            sfb_img = torch.from_numpy(sf_image(sfb)).cuda()
            d = torch.zeros(128, N, device="cuda")
            rc = L.b200q_mxf_probe(a_img.data_ptr(), b_img.data_ptr(), sfa_img.data_ptr(), sfb_img.data_ptr(), d.data_ptr(), N, 1, None, None)
            torch.cuda.synchronize()
            got = d.cpu().numpy().astype(np.float64)
            err = float(np.sum((got - want) ** 2) / np.sum(want ** 2))
            results[(nib_lo_first, data_first)] = err
            print(f"N={N} nibble-low-first={nib_lo_first} data-in-first-8-bytes={data_first}: rc={rc} NMSE vs emulation {err:.3e}", flush=True)
    best = min(results, key=results.get)
    print("best layout:", best, "NMSE", results[best])
    if results[best] < 1e-10:
        # issue rate: 4 MMAs (128 x N x 32 each) per iteration on one SM
        cyc = torch.zeros(1, dtype=torch.int64, device="cuda")
        iters = 2000
        rc = L.b200q_mxf_probe(a_img.data_ptr(), b_img.data_ptr(), sfa_img.data_ptr(), sfb_img.data_ptr(), d.data_ptr(), N, iters, cyc.data_ptr(), None)
        torch.cuda.synchronize()
        c = int(cyc.item())
        per = c / (4 * iters)
        flops = 2 * 128 * N * 32
        print(f"{iters} x 4 MMAs: {c} cycles, {per:.1f} cycles per 128x{N}x32 MMA = {flops / per:.0f} flop/cycle/SM = "
              f"{flops / per * 148 * 1.965e9 / 1e12:.0f} TFLOP/s on 148 SMs at 1965 MHz (dense e2m1 x e4m3)")


if __name__ == "__main__":
    main()
