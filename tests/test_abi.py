"""The C-ABI library loads and exports every symbol include/b200q.h declares (no compute without a GPU)."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared(header):
    src = open(os.path.join(ROOT, "include", header)).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"B200Q_API[^;(]*?\b(b200q_\w+)\s*\(", src)))


def test_header_declares_the_expected_entry_points():
    names = declared("b200q.h")
    for must in ("b200q_mul_mat", "b200q_mul_mat_id", "b200q_get_rows", "b200q_dequantize", "b200q_mul_mat_host"):
        assert must in names


def test_library_exports_every_declared_symbol():
    import b200q
    if not os.path.exists(b200q.LIB_PATH):
        import __graft_entry__ as g
        g.build()
    L = ctypes.CDLL(b200q.LIB_PATH)
    missing = [n for n in declared("b200q.h") if not hasattr(L, n)]
    assert not missing, missing
    assert sorted(b200q.EXPORTS) == [n for n in declared("b200q.h") if n in b200q.EXPORTS]
    assert set(declared("b200q.h")) - set(b200q.EXPORTS) <= set(n for n in declared("b200q.h") if "rowsplit" in n or "graph" in n or "route" in n)


def test_debug_header_matches_the_library():
    """development aids are declared too (include/b200q_debug.h): nothing is exported without a declaration"""
    import subprocess
    import b200q
    L = ctypes.CDLL(b200q.LIB_PATH)
    dbg = declared("b200q_debug.h")
    assert dbg and not [n for n in dbg if not hasattr(L, n)]
    out = subprocess.run(["nm", "-D", "--defined-only", b200q.LIB_PATH], capture_output=True, text=True).stdout
    exported = set(re.findall(r"\b(b200q_\w+)\b", out))
    assert exported <= set(declared("b200q.h")) | set(dbg), exported - set(declared("b200q.h")) - set(dbg)


def test_type_table_matches_ggml_block_sizes():
    import b200q
    L = b200q.lib()
    for t, (be, bb) in b200q.BLCK.items():
        assert L.b200q_blck_size(t) == be
        assert L.b200q_type_size(t) == bb
    assert L.b200q_row_size(b200q.Q4_K, 4096) == 2304
    assert L.b200q_row_size(b200q.MXFP4, 2880) == 1530


def test_compute_fails_loudly_without_a_gpu():
    import torch
    import b200q
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(RuntimeError):
        b200q.dequantize(b200q.QTensor(b200q.Q4_0, (1, 32), None))


def test_plugin_exports_the_ggml_entry_points():
    """libggml-b200.so exports what ggml-backend-reg.cpp:232-276 looks up by name; without a cc 10.0 device
    ggml_backend_score() is 0 (the registry then refuses to load it: there is no CPU fallback to register)"""
    import refs
    import torch
    plugin = os.path.join(ROOT, "llama.cpp-gfx906_b200", "lib", "libggml-b200.so")
    d = refs.ref_dir()
    if d is None or not os.path.exists(plugin):
        pytest.skip("plugin or oracle/_ref not built (needs the reference headers)")
    ctypes.CDLL(os.path.join(d, "libggml-base.so"), mode=ctypes.RTLD_GLOBAL)  # the host process provides ggml-base
    P = ctypes.CDLL(plugin)
    for name in ("ggml_backend_init", "ggml_backend_score", "ggml_backend_b200_reg"):
        assert hasattr(P, name), name
    if not torch.cuda.is_available():
        assert P.ggml_backend_score() == 0
