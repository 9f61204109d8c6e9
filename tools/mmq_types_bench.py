"""prompt-size tensor-core kernel over every weight type: 14336x4096 at n = 512 (cold weights), TFLOP/s by 2 m n k"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tools"))
import quick_bench as qb  # noqa: E402
import refs  # noqa: E402
import torch  # noqa: E402

torch.cuda.set_device(0)
m, k, n = 14336, 4096, 512
for t in refs.WEIGHT_TYPES:
    if k % refs.BLCK[t][0]:
        continue
    us = qb.bench(t, m, k, n, copies=6, iters=60)
    print(f"    -> {2.0 * m * n * k / us / 1e6:7.1f} TFLOP/s", flush=True)
