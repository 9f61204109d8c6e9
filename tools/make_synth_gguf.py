"""Write a synthetic-weight GGUF of Llama-3-8B's architecture (no tokenizer: tokenizer.ggml.model = "no_vocab") for the
reference's llama-bench.  Our own minimal GGUF v3 writer (the format is documented in the reference's ggml/include/gguf.h and
docs; nothing is imported from the reference, so this runs on the GPU box).  Weights are valid random blocks
(tests/refs.random_blocks) tiled over each matrix — throughput does not depend on the values.

python tools/make_synth_gguf.py --out /tmp/synth-8b-q4_k_m.gguf [--layers 32] [--ftype q4_k_m|q5_k_m|q8_0|q4_0]"""
import argparse
import os
import struct
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import refs  # noqa: E402

F32 = 0
ALIGN = 32
T_U32, T_F32, T_STR, T_U64 = 4, 6, 8, 10


def _s(b: bytes) -> bytes:
    return struct.pack("<Q", len(b)) + b


def kv(key: str, typ: int, val) -> bytes:
    out = _s(key.encode()) + struct.pack("<I", typ)
    if typ == T_U32:
        out += struct.pack("<I", val)
    elif typ == T_U64:
        out += struct.pack("<Q", val)
    elif typ == T_F32:
        out += struct.pack("<f", val)
    elif typ == T_STR:
        out += _s(val.encode())
    else:
        raise ValueError(typ)
    return out


def use_more_bits(i, n):  # llama-quant.cpp: which layers get the wider type in the *_M mixtures
    return i < n // 8 or i >= 7 * n // 8 or (i - n // 8) % 3 == 2


def write_gguf(path, kvs, tensors):
    """tensors: (name, ne in ggml order, type); data = valid random blocks tiled / f32 noise"""
    def nbytes(ne, t):
        rows = int(np.prod(ne[1:])) if len(ne) > 1 else 1
        return rows * (ne[0] * 4 if t == F32 else refs.row_size(t, ne[0]))

    head = struct.pack("<IIQQ", 0x46554747, 3, len(tensors), len(kvs)) + b"".join(kvs)
    off, infos = 0, b""
    for name, ne, t in tensors:
        infos += _s(name.encode()) + struct.pack("<I", len(ne)) + b"".join(struct.pack("<Q", d) for d in ne)
        infos += struct.pack("<IQ", t, off)
        off += (nbytes(ne, t) + ALIGN - 1) // ALIGN * ALIGN
    blocks = {}
    with open(path, "wb") as f:
        f.write(head + infos)
        f.write(b"\0" * (-f.tell() % ALIGN))
        for name, ne, t in tensors:
            rows = int(np.prod(ne[1:])) if len(ne) > 1 else 1
            if t == F32:
                scale = 1.0 if name.endswith("norm.weight") else 0.0
                data = (scale + 0.01 * np.random.default_rng(len(name)).standard_normal(ne[0] * rows)).astype(np.float32).tobytes()
            else:
                key = (t, ne[0])
                if key not in blocks:
                    blocks[key] = refs.random_blocks(t, 256, ne[0], seed=17 + t)
                data = np.tile(blocks[key], ((rows + 255) // 256, 1))[:rows].tobytes()
            assert len(data) == nbytes(ne, t), (name, len(data), nbytes(ne, t))
            f.write(data)
            f.write(b"\0" * (-len(data) % ALIGN))
    print(f"{path}: {len(tensors)} tensors, {os.path.getsize(path) / 1e9:.2f} GB")


def gpt_oss(a):
    """gpt-oss-20B's architecture (src/llama-model.cpp LLM_ARCH_OPENAI_MOE): MXFP4 experts, q8_0 attention, f32 biases / sinks / router"""
    n_embd, n_head, n_head_kv, hd, n_expert, n_used, n_ff = 2880, 64, 8, 64, 32, 4, 2880
    layers = a.layers if a.layers != 32 else 24
    vocab = a.vocab or 201088
    Q8, MX = refs.Q8_0, refs.MXFP4
    tensors = [("token_embd.weight", (n_embd, vocab), Q8)]
    for i in range(layers):
        p = f"blk.{i}."
        tensors += [(p + "attn_norm.weight", (n_embd,), F32), (p + "post_attention_norm.weight", (n_embd,), F32),
                    (p + "attn_q.weight", (n_embd, n_head * hd), Q8), (p + "attn_q.bias", (n_head * hd,), F32),
                    (p + "attn_k.weight", (n_embd, n_head_kv * hd), Q8), (p + "attn_k.bias", (n_head_kv * hd,), F32),
                    (p + "attn_v.weight", (n_embd, n_head_kv * hd), Q8), (p + "attn_v.bias", (n_head_kv * hd,), F32),
                    (p + "attn_output.weight", (n_head * hd, n_embd), Q8), (p + "attn_output.bias", (n_embd,), F32),
                    (p + "attn_sinks.weight", (n_head,), F32),
                    (p + "ffn_gate_inp.weight", (n_embd, n_expert), F32), (p + "ffn_gate_inp.bias", (n_expert,), F32),
                    (p + "ffn_gate_exps.weight", (n_embd, n_ff, n_expert), MX), (p + "ffn_gate_exps.bias", (n_ff, n_expert), F32),
                    (p + "ffn_up_exps.weight", (n_embd, n_ff, n_expert), MX), (p + "ffn_up_exps.bias", (n_ff, n_expert), F32),
                    (p + "ffn_down_exps.weight", (n_ff, n_embd, n_expert), MX), (p + "ffn_down_exps.bias", (n_embd, n_expert), F32)]
    tensors += [("output_norm.weight", (n_embd,), F32), ("output.weight", (n_embd, vocab), Q8)]
    A = "gpt-oss."
    kvs = [kv("general.architecture", T_STR, "gpt-oss"), kv("general.name", T_STR, "synthetic gpt-oss-20b-shaped mxfp4"),
           kv("general.alignment", T_U32, ALIGN), kv(A + "context_length", T_U32, 8192), kv(A + "embedding_length", T_U32, n_embd),
           kv(A + "block_count", T_U32, layers), kv(A + "feed_forward_length", T_U32, n_ff), kv(A + "expert_feed_forward_length", T_U32, n_ff),
           kv(A + "attention.head_count", T_U32, n_head), kv(A + "attention.head_count_kv", T_U32, n_head_kv),
           kv(A + "attention.key_length", T_U32, hd), kv(A + "attention.value_length", T_U32, hd),
           kv(A + "attention.layer_norm_rms_epsilon", T_F32, 1e-5), kv(A + "attention.sliding_window", T_U32, 128),
           kv(A + "rope.dimension_count", T_U32, hd), kv(A + "rope.freq_base", T_F32, 150000.0),
           kv(A + "expert_count", T_U32, n_expert), kv(A + "expert_used_count", T_U32, n_used), kv(A + "vocab_size", T_U32, vocab),
           kv("tokenizer.ggml.model", T_STR, "no_vocab")]
    write_gguf(a.out, kvs, tensors)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", required=True)
    ap.add_argument("--layers", type=int, default=32)
    ap.add_argument("--ftype", default="q4_k_m")
    ap.add_argument("--vocab", type=int, default=0)
    ap.add_argument("--arch", default="llama", choices=["llama", "gpt-oss"])
    a = ap.parse_args()
    if a.arch == "gpt-oss":
        return gpt_oss(a)
    a.vocab = a.vocab or 128256
    n_embd, n_ff, n_head, n_head_kv, hd = 4096, 14336, 32, 8, 128
    base = {"q4_k_m": refs.Q4_K, "q5_k_m": refs.Q5_K, "q8_0": refs.Q8_0, "q4_0": refs.Q4_0}[a.ftype]
    wide = refs.Q6_K if a.ftype in ("q4_k_m", "q5_k_m") else base
    tensors = []  # (name, ne (ggml order), type)
    tensors.append(("token_embd.weight", (n_embd, a.vocab), base))
    for i in range(a.layers):
        more = use_more_bits(i, a.layers)
        p = f"blk.{i}."
        tensors += [(p + "attn_norm.weight", (n_embd,), F32),
                    (p + "attn_q.weight", (n_embd, n_embd), base),
                    (p + "attn_k.weight", (n_embd, hd * n_head_kv), base),
                    (p + "attn_v.weight", (n_embd, hd * n_head_kv), wide if more else base),
                    (p + "attn_output.weight", (n_embd, n_embd), base),
                    (p + "ffn_norm.weight", (n_embd,), F32),
                    (p + "ffn_gate.weight", (n_embd, n_ff), base),
                    (p + "ffn_up.weight", (n_embd, n_ff), base),
                    (p + "ffn_down.weight", (n_ff, n_embd), wide if more else base)]
    tensors += [("output_norm.weight", (n_embd,), F32), ("output.weight", (n_embd, a.vocab), wide)]

    def nbytes(ne, t):
        rows = int(np.prod(ne[1:])) if len(ne) > 1 else 1
        return rows * (ne[0] * 4 if t == F32 else refs.row_size(t, ne[0]))

    kvs = [kv("general.architecture", T_STR, "llama"), kv("general.name", T_STR, f"synthetic llama3-8b-shaped {a.ftype}"),
           kv("general.alignment", T_U32, ALIGN),
           kv("llama.context_length", T_U32, 8192), kv("llama.embedding_length", T_U32, n_embd),
           kv("llama.block_count", T_U32, a.layers), kv("llama.feed_forward_length", T_U32, n_ff),
           kv("llama.attention.head_count", T_U32, n_head), kv("llama.attention.head_count_kv", T_U32, n_head_kv),
           kv("llama.attention.layer_norm_rms_epsilon", T_F32, 1e-5), kv("llama.rope.dimension_count", T_U32, hd),
           kv("llama.rope.freq_base", T_F32, 500000.0), kv("llama.vocab_size", T_U32, a.vocab),
           kv("tokenizer.ggml.model", T_STR, "no_vocab")]
    head = struct.pack("<IIQQ", 0x46554747, 3, len(tensors), len(kvs)) + b"".join(kvs)
    off = 0
    infos = b""
    for name, ne, t in tensors:
        infos += _s(name.encode()) + struct.pack("<I", len(ne)) + b"".join(struct.pack("<Q", d) for d in ne)
        infos += struct.pack("<IQ", t, off)
        off += (nbytes(ne, t) + ALIGN - 1) // ALIGN * ALIGN
    blocks = {}  # (type, k) -> 256 valid random rows
    with open(a.out, "wb") as f:
        f.write(head + infos)
        f.write(b"\0" * (-f.tell() % ALIGN))
        for name, ne, t in tensors:
            if t == F32:
                data = (1.0 + 0.01 * np.random.default_rng(len(name)).standard_normal(ne[0])).astype(np.float32).tobytes()
            else:
                key = (t, ne[0])
                if key not in blocks:
                    blocks[key] = refs.random_blocks(t, 256, ne[0], seed=17 + t)
                rows = ne[1]
                data = np.tile(blocks[key], ((rows + 255) // 256, 1))[:rows].tobytes()
            assert len(data) == nbytes(ne, t), (name, len(data), nbytes(ne, t))
            f.write(data)
            f.write(b"\0" * (-len(data) % ALIGN))
    print(f"{a.out}: {len(tensors)} tensors, {os.path.getsize(a.out) / 1e9:.2f} GB")


if __name__ == "__main__":
    main()
