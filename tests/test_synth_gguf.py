"""tools/make_synth_gguf.py writes the synthetic-weight GGUFs the llama-bench legs run on (our own GGUF v3 writer).  Here the
reference's own parser (gguf_init_from_file of the unmodified ggml-base in oracle/_ref, test infrastructure) reads a small file
back: tensor count, names, types, sizes and offsets must be what the writer meant."""
import ctypes as C
import os
import subprocess
import sys

import pytest

import refs

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "oracle", "_ref", "libggml-base.so")


class _Params(C.Structure):
    _fields_ = [("no_alloc", C.c_bool), ("ctx", C.c_void_p)]


@pytest.mark.skipif(not os.path.exists(LIB), reason="oracle/_ref not built (make -C oracle)")
@pytest.mark.parametrize("ftype,base,wide", [("q4_k_m", refs.Q4_K, refs.Q6_K), ("q5_k_m", refs.Q5_K, refs.Q6_K), ("q8_0", refs.Q8_0, refs.Q8_0)])
def test_reference_parser_reads_the_synthetic_gguf(tmp_path, ftype, base, wide):
    path = str(tmp_path / "synth.gguf")
    subprocess.run([sys.executable, os.path.join(ROOT, "tools", "make_synth_gguf.py"), "--out", path, "--layers", "1", "--vocab", "512",
                    "--ftype", ftype], check=True, capture_output=True, timeout=300)
    g = C.CDLL(LIB)
    g.gguf_init_from_file.restype = C.c_void_p
    g.gguf_init_from_file.argtypes = [C.c_char_p, _Params]
    for fn, rt in (("gguf_get_n_tensors", C.c_int64), ("gguf_get_n_kv", C.c_int64), ("gguf_get_tensor_name", C.c_char_p),
                   ("gguf_get_tensor_type", C.c_int), ("gguf_get_tensor_offset", C.c_size_t), ("gguf_get_tensor_size", C.c_size_t),
                   ("gguf_get_data_offset", C.c_size_t), ("gguf_find_key", C.c_int64), ("gguf_get_val_str", C.c_char_p)):
        getattr(g, fn).restype = rt
    for fn in ("gguf_get_tensor_name", "gguf_get_tensor_type", "gguf_get_tensor_offset", "gguf_get_tensor_size"):
        getattr(g, fn).argtypes = [C.c_void_p, C.c_int64]
    g.gguf_get_n_tensors.argtypes = g.gguf_get_n_kv.argtypes = g.gguf_get_data_offset.argtypes = g.gguf_free.argtypes = [C.c_void_p]
    g.gguf_find_key.argtypes = [C.c_void_p, C.c_char_p]
    g.gguf_get_val_str.argtypes = [C.c_void_p, C.c_int64]
    ctx = g.gguf_init_from_file(path.encode(), _Params(True, None))
    assert ctx, "the reference's parser rejected the file"
    try:
        assert g.gguf_get_n_tensors(ctx) == 1 + 9 + 2  # token_embd, one layer, output_norm + output
        assert g.gguf_get_val_str(ctx, g.gguf_find_key(ctx, b"general.architecture")) == b"llama"
        assert g.gguf_get_val_str(ctx, g.gguf_find_key(ctx, b"tokenizer.ggml.model")) == b"no_vocab"
        seen, end = {}, 0
        for i in range(12):
            name = g.gguf_get_tensor_name(ctx, i).decode()
            seen[name] = (g.gguf_get_tensor_type(ctx, i), g.gguf_get_tensor_size(ctx, i))
            assert g.gguf_get_tensor_offset(ctx, i) == end and end % 32 == 0, name  # dense, 32-byte aligned, in order
            end += (seen[name][1] + 31) // 32 * 32
        assert g.gguf_get_data_offset(ctx) + end == os.path.getsize(path)
        assert seen["blk.0.attn_q.weight"] == (base, 4096 * refs.row_size(base, 4096))
        assert seen["blk.0.ffn_down.weight"] == (wide, 4096 * refs.row_size(wide, 14336))  # layer 0 of 1 takes the wider type in *_M
        assert seen["output.weight"] == (wide, 512 * refs.row_size(wide, 4096))
        assert seen["blk.0.attn_norm.weight"] == (0, 4096 * 4)
    finally:
        g.gguf_free(ctx)
