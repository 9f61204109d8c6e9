"""Host logic of the row-split mode on CPU: the partition rule (reference get_row_split, ggml-cuda.cu:760-786) and the
N > 1 gather path with world_size 2 over gloo.  The per-rank compute is INJECTED here (the oracle on a numpy shard):
this only checks sharding / layout / collective plumbing — the product path uses the CUDA kernels + NCCL and is
checked bit-for-bit on GPUs in test_gpu_rowsplit.py."""
import os
import socket

import numpy as np
import pytest

import refs


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def test_row_split_partition_rule():
    from b200q.rowsplit import row_split
    assert row_split(4096, 1) == [(0, 4096)]
    assert row_split(4096, 2) == [(0, 2048), (2048, 2048)]
    assert row_split(28672, 8) == [(3584 * r, 3584) for r in range(8)]
    # rounded DOWN to the 128-row granule, last rank takes the remainder
    assert row_split(300, 2) == [(0, 128), (128, 172)]
    assert row_split(1000, 4) == [(0, 128), (128, 256), (384, 256), (640, 360)]
    # uneven tensor_split, normalised
    parts = row_split(4096, 2, tensor_split=[3, 1])
    assert parts == [(0, 3072), (3072, 1024)]
    # tiny matrices: everything lands on the last rank, empty shards are legal
    assert row_split(16, 4) == [(0, 0), (0, 0), (0, 0), (0, 16)]
    for m in (1, 127, 128, 129, 4096, 128256):
        for w in (1, 2, 3, 4, 8):
            parts = row_split(m, w)
            assert parts[0][0] == 0 and sum(n for _, n in parts) == m
            for (a, n), (b, _) in zip(parts, parts[1:]):
                assert a + n == b and a % 128 == 0
    with pytest.raises(ValueError):
        row_split(10, 2, tensor_split=[1.0])


def test_shard_bytes_are_row_slices():
    from b200q.rowsplit import shard_bytes
    t, m, k = refs.Q4_K, 300, 512
    w = refs.make_weights(t, m, k, seed=3)
    s = shard_bytes(t, m, k, w, 128, 172)
    assert s.shape == (172, refs.row_size(t, k))
    assert s.tobytes() == np.ascontiguousarray(w).view(np.uint8).reshape(m, -1)[128:].tobytes()


def _worker(rank, world, port, t, m, k, n, q):
    import torch
    import torch.distributed as dist
    from b200q import QTensor
    from b200q.rowsplit import mul_mat_row_split, upload_row_split
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        w = refs.make_weights(t, m, k, seed=11)
        act = refs.uniform_weights(n, k, seed=12)

        def fake_upload(qtype, shape, host_bytes):
            return QTensor(qtype, tuple(shape), np.ascontiguousarray(host_bytes))

        def oracle_compute(shard, a, out):   # test infrastructure only
            out.copy_(torch.from_numpy(refs.orc_mul_mat(shard.qtype, shard.data, a.numpy())))

        ws = upload_row_split(t, (m, k), w, rank, world, uploader=fake_upload)
        got = mul_mat_row_split(ws, torch.from_numpy(act), compute=oracle_compute).numpy()
        want = refs.orc_mul_mat(t, w, act)
        q.put((rank, got.tobytes() == want.tobytes(), ws.parts))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("t,m,k,n", [(refs.Q4_K, 512, 256, 1), (refs.Q8_0, 300, 256, 3), (refs.Q4_0, 64, 128, 2)],
                         ids=["equal-n1", "ragged-n3", "empty-shard"])
def test_row_split_gather_gloo_world2(t, m, k, n):
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, t, m, k, n, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in procs]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert all(ok for _, ok, _ in res), res
