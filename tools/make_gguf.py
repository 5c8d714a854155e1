#!/usr/bin/env python3
"""Random-init TinyLlama-1.1B-shaped Q4_K_M GGUF for the reference's own llama-bench (README.md:189-193).

Standalone GGUF v3 writer (numpy only: it has to run on the GPU box, where neither /root/reference nor gguf-py exist).
Format: L/ggml/src/gguf.cpp (header, KV section, tensor infos, 32-byte aligned data).  Tensor types follow what
`llama-quantize ... Q4_K_M` produces for this architecture (L/src/llama-quant.cpp:185-187, 228-230, 302-303, 358-364):
everything Q4_K, except attn_v / ffn_down of layers {0,1,4,7,10,13,16,19,20,21} and output.weight in Q6_K; norms F32.
The vocabulary is `no_vocab` with llama.vocab_size set (L/src/llama-vocab.cpp:1679-1697): llama-bench feeds random token ids
(L/tools/llama-bench/llama-bench.cpp:1922-1939) and never detokenizes.

Weights are random but SANE block bytes: fp16 d / dmin around 1e-3, random 6-bit scales / mins, random nibbles — the
activations of a 22-layer pass stay finite, which keeps the CPU-side ops (softmax, norms) on their normal paths.
"""
import argparse
import struct

import numpy as np

F32, Q4_K, Q6_K = 0, 12, 14
T_U32, T_F32, T_STR = 4, 6, 8
ALIGN = 32

TINYLLAMA = dict(n_layer=22, n_embd=2048, n_head=32, n_head_kv=4, n_ff=5632, n_vocab=32000, n_ctx=2048)
Q6_LAYERS = {0, 1, 4, 7, 10, 13, 16, 19, 20, 21}      # use_more_bits(i, 22): i < 3 of 8ths / last 8th / every third between


def random_q4_K(rng, rows, k):
    nb = k // 256
    blk = np.empty((rows * nb, 144), np.uint8)
    hdr = np.empty((rows * nb, 2), np.float16)
    hdr[:, 0] = rng.uniform(5e-4, 2e-3, rows * nb)           # d
    hdr[:, 1] = rng.uniform(5e-4, 2e-3, rows * nb)           # dmin
    blk[:, 0:4] = hdr.view(np.uint8)
    blk[:, 4:] = rng.integers(0, 256, (rows * nb, 140), dtype=np.uint8)
    return blk.reshape(-1)


def random_q6_K(rng, rows, k):
    nb = k // 256
    blk = np.empty((rows * nb, 210), np.uint8)
    blk[:, :192] = rng.integers(0, 256, (rows * nb, 192), dtype=np.uint8)          # ql, qh
    blk[:, 192:208] = rng.integers(-64, 64, (rows * nb, 16), dtype=np.int8).view(np.uint8)   # scales
    blk[:, 208:210] = rng.uniform(2e-5, 8e-5, rows * nb).astype(np.float16).reshape(-1, 1).view(np.uint8)
    return blk.reshape(-1)


def gstr(s):
    b = s.encode()
    return struct.pack("<Q", len(b)) + b


def kv(key, typ, val):
    out = gstr(key) + struct.pack("<I", typ)
    if typ == T_U32:
        out += struct.pack("<I", val)
    elif typ == T_F32:
        out += struct.pack("<f", val)
    else:
        out += gstr(val)
    return out


def tensor_list(c):
    """(name, ne (fastest first, as ggml), type) in file order"""
    E, FF, V = c["n_embd"], c["n_ff"], c["n_vocab"]
    kvd = E // c["n_head"] * c["n_head_kv"]
    ts = [("token_embd.weight", (E, V), Q4_K), ("output_norm.weight", (E,), F32), ("output.weight", (E, V), Q6_K)]
    for i in range(c["n_layer"]):
        six = Q6_K if i in Q6_LAYERS else Q4_K
        ts += [(f"blk.{i}.attn_norm.weight", (E,), F32), (f"blk.{i}.attn_q.weight", (E, E), Q4_K),
               (f"blk.{i}.attn_k.weight", (E, kvd), Q4_K), (f"blk.{i}.attn_v.weight", (E, kvd), six),
               (f"blk.{i}.attn_output.weight", (E, E), Q4_K), (f"blk.{i}.ffn_norm.weight", (E,), F32),
               (f"blk.{i}.ffn_gate.weight", (E, FF), Q4_K), (f"blk.{i}.ffn_up.weight", (E, FF), Q4_K),
               (f"blk.{i}.ffn_down.weight", (FF, E), six)]
    return ts


def nbytes(ne, typ):
    n = int(np.prod(ne))
    return n * 4 if typ == F32 else n // 256 * (144 if typ == Q4_K else 210)


def write(path, c=TINYLLAMA, seed=1234, all_q4k=False):
    rng = np.random.default_rng(seed)
    ts = tensor_list(c)
    if all_q4k:
        ts = [(n, ne, Q4_K if t == Q6_K else t) for n, ne, t in ts]
    kvs = [kv("general.architecture", T_STR, "llama"), kv("general.name", T_STR, "tinyllama-1.1b-random-init"),
           kv("general.file_type", T_U32, 15),                       # LLAMA_FTYPE_MOSTLY_Q4_K_M
           kv("llama.context_length", T_U32, c["n_ctx"]), kv("llama.embedding_length", T_U32, c["n_embd"]),
           kv("llama.block_count", T_U32, c["n_layer"]), kv("llama.feed_forward_length", T_U32, c["n_ff"]),
           kv("llama.rope.dimension_count", T_U32, c["n_embd"] // c["n_head"]),
           kv("llama.attention.head_count", T_U32, c["n_head"]), kv("llama.attention.head_count_kv", T_U32, c["n_head_kv"]),
           kv("llama.attention.layer_norm_rms_epsilon", T_F32, 1e-5), kv("llama.rope.freq_base", T_F32, 10000.0),
           kv("llama.vocab_size", T_U32, c["n_vocab"]), kv("tokenizer.ggml.model", T_STR, "no_vocab")]
    head = b"GGUF" + struct.pack("<IQQ", 3, len(ts), len(kvs)) + b"".join(kvs)
    infos, off = b"", 0
    for name, ne, typ in ts:
        infos += gstr(name) + struct.pack("<I", len(ne)) + b"".join(struct.pack("<Q", d) for d in ne) + struct.pack("<IQ", typ, off)
        off += (nbytes(ne, typ) + ALIGN - 1) // ALIGN * ALIGN
    with open(path, "wb") as f:
        f.write(head + infos)
        f.write(b"\0" * (-f.tell() % ALIGN))
        for name, ne, typ in ts:
            if typ == F32:
                data = np.ones(ne[0], np.float32).view(np.uint8)
            else:
                rows = int(np.prod(ne[1:]))
                data = (random_q4_K if typ == Q4_K else random_q6_K)(rng, rows, ne[0])
            assert data.size == nbytes(ne, typ), name
            f.write(data.tobytes())
            f.write(b"\0" * (-data.size % ALIGN))
    q4 = sum(nbytes(ne, t) for n, ne, t in ts if t == Q4_K and n != "token_embd.weight")
    q6 = sum(nbytes(ne, t) for n, ne, t in ts if t == Q6_K)
    return {"tensors": len(ts), "q4k_matmul_bytes": q4, "q6k_matmul_bytes": q6, "file_bytes": off}


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("out")
    ap.add_argument("--seed", type=int, default=1234)
    ap.add_argument("--all-q4k", action="store_true", help="every quantized tensor Q4_K (the 545 MB all-Q4_K variant of SURVEY.md §8d)")
    ap.add_argument("--layers", type=int, default=TINYLLAMA["n_layer"])
    a = ap.parse_args()
    cfg = dict(TINYLLAMA, n_layer=a.layers)
    print(write(a.out, cfg, a.seed, a.all_q4k))
