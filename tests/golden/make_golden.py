"""Generate the committed golden vectors from the UNMODIFIED reference (oracle/_ref, built by oracle/Makefile
from /root/reference).  Run in the build container:  python tests/golden/make_golden.py

For every weight type the fixture holds, all produced by the reference's own code:
  wq      quantized rows  (ggml_quantize_chunk, imatrix = NULL)        of seeded U(-1,1) weights [M, K]
  wq_cos  quantized rows of the reference's deterministic fixture 0.1 + 2 cos(i + offset)
          (tests/test-quantize-fns.cpp:31-35)
  deq / deq_cos   type_traits.to_float of those rows (the dequantize / GET_ROWS oracle)
  act     f32 activations [N, K], seeded U(-1,1)
  actq    act quantized to the CPU vec_dot type (type_traits_cpu.from_float)
  dots    ggml_vec_dot(wq[i], actq[j])  [N, M]       (SIMD variant of this build)
  mm      ggml_mul_mat on the CPU backend, 4 threads  [N, M]
  mmid    ggml_mul_mat_id, experts = wq reshaped [E, M/E, K]
"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.join(os.path.dirname(__file__), ".."))
import refs  # noqa: E402

M, K, N, E = 8, 512, 3, 4


def main():
    R = refs.ref()
    assert R is not None, "oracle/_ref not built (make -C oracle ref)"
    print("reference CPU features:", R.ref_cpu_features().decode(), "dir:", refs.ref_dir())
    out = {}
    for t in refs.WEIGHT_TYPES:
        name = refs.NAMES[t]
        w = refs.uniform_weights(M, K, 1234 + t)
        wcos = (0.1 + 2.0 * np.cos(np.arange(M * K, dtype=np.float32) + np.float32(t))).astype(np.float32).reshape(M, K)
        rs = refs.row_size(t, K)
        for tag, src in (("", w), ("_cos", wcos)):
            q = np.zeros((M, rs), dtype=np.uint8)
            R.ref_quantize(t, refs._p(np.ascontiguousarray(src)), refs._p(q), M, K)
            d = np.empty((M, K), dtype=np.float32)
            assert R.ref_dequantize(t, refs._p(q), refs._p(d), M * K) == 0
            out[f"{name}/wq{tag}"] = q
            out[f"{name}/deq{tag}"] = d
        q = out[f"{name}/wq"]
        act = refs.uniform_weights(N, K, 4321 + t)
        vt = R.ref_vec_dot_type(t)
        aq = np.zeros((N, refs.row_size(vt, K)), dtype=np.uint8)
        for j in range(N):
            assert R.ref_act_quantize(t, refs._p(act[j]), refs._p(aq[j]), K) == vt
        dots = np.zeros((N, M), dtype=np.float32)
        for j in range(N):
            for i in range(M):
                o = np.zeros(1, dtype=np.float32)
                assert R.ref_vec_dot(t, K, refs._p(q[i]), refs._p(aq[j]), refs._p(o)) == 0
                dots[j, i] = o[0]
        mm, _ = refs.ref_mul_mat(t, q, act, threads=4)
        ids = np.array([[0, 2, 9], [3, 1, 9], [1, 0, 9]], dtype=np.int32)  # stride 3, n_used 2 (a view, as in the harness)
        b = refs.uniform_weights(N * 2, K, 999 + t).reshape(N, 2, K)
        mmid = np.empty((N, 2, M // E), dtype=np.float32)
        secs = R.ref_mul_mat_id(t, refs._p(q), refs._p(b), refs._p(ids), refs._p(mmid), M // E, K, E, 2, N, 2, 3, 4, 1)
        assert secs >= 0
        out[f"{name}/act"] = act
        out[f"{name}/actq"] = aq
        out[f"{name}/dots"] = dots
        out[f"{name}/mm"] = mm
        out[f"{name}/ids"] = ids
        out[f"{name}/b"] = b
        out[f"{name}/mmid"] = mmid
        out[f"{name}/vec_dot_type"] = np.array([vt], dtype=np.int32)
    path = os.path.join(os.path.dirname(__file__), "ref_vectors.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
