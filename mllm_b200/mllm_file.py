""".mllm weight container (reader / writer) and the model configurations of the BASELINE configs.

File format (mllm/ParamLoader.cpp:15-31,157-286; writer: tools/convertor/converter.py `Writer`):
    int32 magic = 20012 | uint64 index_len | index entries | tensor bytes
    entry = int32 name_len, name bytes, uint64 size, uint64 offset (from file start), int32 dtype
Tensor bytes are the native ggml-style blocks (mllm/DataType.hpp) and are uploaded as they are.
"""
import os
import struct

import numpy as np

MAGIC = 20012
F32, F16, Q4_0, Q8_0 = 0, 1, 2, 8
ROPE_LLAMA, ROPE_HF = 2, 4


class Names:
    """LLaMANameConfig::init (models/llama/configuration_llama.hpp:20-60)."""

    def __init__(self, rope_type):
        if rope_type == ROPE_LLAMA:
            self.blk, self.attn, self.ffn = "layers.", "attention.", "feed_forward."
            self.q, self.k, self.v, self.o = "wq", "wk", "wv", "wo"
            self.gate, self.up, self.down = "w1", "w3", "w2"
            self.attn_norm, self.ffn_norm = "attention_norm", "ffn_norm"
            self.embed, self.norm, self.lm_head = "tok_embeddings.weight", "norm.weight", "output.weight"
        else:
            self.blk, self.attn, self.ffn = "model.layers.", "self_attn.", "mlp."
            self.q, self.k, self.v, self.o = "q_proj", "k_proj", "v_proj", "o_proj"
            self.gate, self.up, self.down = "gate_proj", "up_proj", "down_proj"
            self.attn_norm, self.ffn_norm = "input_layernorm", "post_attention_layernorm"
            self.embed, self.norm, self.lm_head = "model.embed_tokens.weight", "model.norm.weight", "lm_head.weight"

    def layer(self, L, role):
        b = f"{self.blk}{L}."
        a, f = b + self.attn, b + self.ffn
        return {
            "attn_norm": b + self.attn_norm + ".weight", "ffn_norm": b + self.ffn_norm + ".weight",
            "wq": a + self.q + ".weight", "wk": a + self.k + ".weight", "wv": a + self.v + ".weight", "wo": a + self.o + ".weight",
            "bq": a + self.q + ".bias", "bk": a + self.k + ".bias", "bv": a + self.v + ".bias",
            "w_gate": f + self.gate + ".weight", "w_up": f + self.up + ".weight", "w_down": f + self.down + ".weight",
        }[role]


# shapes from the reference's config classes
CONFIGS = {
    "tinyllama-1.1b": dict(vocab=32000, hidden=2048, layers=22, heads=32, kv_heads=4, ffn=5632, rope_type=ROPE_HF, rope_theta=10000.0, rms_eps=1e-5, qkv_bias=0),
    "llama2-7b": dict(vocab=32000, hidden=4096, layers=32, heads=32, kv_heads=32, ffn=11008, rope_type=ROPE_LLAMA, rope_theta=10000.0, rms_eps=1e-6, qkv_bias=0),
    "qwen1.5-1.8b": dict(vocab=151936, hidden=2048, layers=24, heads=16, kv_heads=16, ffn=5504, rope_type=ROPE_HF, rope_theta=1000000.0, rms_eps=1e-6, qkv_bias=1),
    # small graphs with the same structure, for parity tests the CPU oracle finishes in seconds
    "toy-llama": dict(vocab=512, hidden=256, layers=2, heads=4, kv_heads=4, ffn=512, rope_type=ROPE_LLAMA, rope_theta=10000.0, rms_eps=1e-6, qkv_bias=0),
    "toy-d128": dict(vocab=320, hidden=256, layers=2, heads=2, kv_heads=1, ffn=512, rope_type=ROPE_LLAMA, rope_theta=10000.0, rms_eps=1e-6, qkv_bias=0),
    "toy-gqa": dict(vocab=384, hidden=256, layers=3, heads=8, kv_heads=2, ffn=768, rope_type=ROPE_HF, rope_theta=10000.0, rms_eps=1e-5, qkv_bias=1),
}


def config(name, cache_limit=1024, max_pos=None):
    c = dict(CONFIGS[name])
    c["name"] = name
    c["head_dim"] = c["hidden"] // c["heads"]
    c["cache_limit"] = cache_limit
    c["pos_max"] = max_pos or cache_limit
    return c


def tensor_list(cfg):
    """[(name, rows, cols, dtype)] of every parameter, in file order."""
    n = Names(cfg["rope_type"])
    H, D = cfg["hidden"], cfg["head_dim"]
    out = [(n.embed, cfg["vocab"], H, Q4_0)]
    for L in range(cfg["layers"]):
        out.append((n.layer(L, "attn_norm"), 1, H, F32))
        for role, rows in (("wq", cfg["heads"] * D), ("wk", cfg["kv_heads"] * D), ("wv", cfg["kv_heads"] * D)):
            out.append((n.layer(L, role), rows, H, Q4_0))
            if cfg["qkv_bias"]:
                out.append((n.layer(L, "b" + role[1]), 1, rows, F32))
        out.append((n.layer(L, "wo"), H, cfg["heads"] * D, Q4_0))
        out.append((n.layer(L, "ffn_norm"), 1, H, F32))
        out.append((n.layer(L, "w_gate"), cfg["ffn"], H, Q4_0))
        out.append((n.layer(L, "w_up"), cfg["ffn"], H, Q4_0))
        out.append((n.layer(L, "w_down"), H, cfg["ffn"], Q4_0))
    out.append((n.norm, 1, H, F32))
    out.append((n.lm_head, cfg["vocab"], H, Q4_0))
    return out


def nbytes(rows, cols, dtype):
    return rows * cols * 4 if dtype == F32 else rows * (cols // 32) * 18


def quantize_q4_0_np(w):
    """Vectorised numpy twin of quantize_row_q4_0_reference (ggml/QuantizeQ4.cpp:33-68).
    Used only to MAKE random-init model files; its bytes are pinned against the oracle in tests."""
    w = np.ascontiguousarray(w, dtype=np.float32)
    rows, cols = w.shape
    b = w.reshape(rows, cols // 32, 32)
    idx = np.abs(b).argmax(axis=2)  # first maximum, as the reference's strict '<' keeps
    mx = np.take_along_axis(b, idx[..., None], axis=2)[..., 0]
    d = (mx / np.float32(-8)).astype(np.float32)
    with np.errstate(divide="ignore"):
        idv = np.where(d != 0, np.float32(1.0) / d, np.float32(0)).astype(np.float32)
    x = (b * idv[..., None]).astype(np.float32) + np.float32(8.5)
    q = np.minimum(x.astype(np.int8), 15).astype(np.uint8)
    out = np.empty((rows, cols // 32, 18), dtype=np.uint8)
    out[..., 0:2] = d.astype(np.float16).view(np.uint8).reshape(rows, cols // 32, 2)
    out[..., 2:] = q[..., :16] | (q[..., 16:] << 4)
    return out.reshape(rows, cols // 32 * 18)


def random_tensors(cfg, seed=1234, quantize=None):
    """Random-init weights as SURVEY 8(d) prescribes: W ~ N(0, 0.02^2) -> Q4_0, norms 1 + N(0, 0.02^2), biases N(0, 0.02^2)."""
    quantize = quantize or quantize_q4_0_np
    rng = np.random.default_rng(seed)
    t = {}
    for name, rows, cols, dt in tensor_list(cfg):
        if dt == F32:
            base = 1.0 if "norm" in name else 0.0
            t[name] = (base + 0.02 * rng.standard_normal(cols)).astype(np.float32)
        else:
            t[name] = quantize((0.02 * rng.standard_normal((rows, cols))).astype(np.float32))
    return t


def write_mllm(path, tensors, dtypes):
    """tensors: ordered dict name -> contiguous numpy array (raw bytes are written); dtypes: name -> DataType."""
    names = list(tensors)
    index_len = sum(4 + len(n.encode()) + 8 + 8 + 4 for n in names)
    off = 4 + 8 + index_len
    with open(path, "wb") as f:
        f.write(struct.pack("<i", MAGIC))
        f.write(struct.pack("<Q", index_len))
        for n in names:
            nb = tensors[n].nbytes
            e = n.encode()
            f.write(struct.pack("<i", len(e)) + e + struct.pack("<QQi", nb, off, dtypes[n]))
            off += nb
        for n in names:
            f.write(np.ascontiguousarray(tensors[n]).tobytes() if tensors[n].nbytes < (1 << 28) else memoryview(np.ascontiguousarray(tensors[n])))
    return path


def write_model(path, cfg, tensors):
    dt = {name: d for name, _, _, d in tensor_list(cfg)}
    return write_mllm(path, {n: tensors[n] for n in dt}, dt)


def read_mllm(path):
    """-> (dict name -> np.memmap view (uint8 for block types, float32 for F32), dict name -> dtype)."""
    mm = np.memmap(path, dtype=np.uint8, mode="r")
    magic, = struct.unpack_from("<i", mm, 0)
    if magic != MAGIC:
        raise ValueError("magic number error")
    index_len, = struct.unpack_from("<Q", mm, 4)
    pos, end = 12, 12 + index_len
    tensors, dtypes = {}, {}
    while pos < end:
        ln, = struct.unpack_from("<i", mm, pos)
        name = bytes(mm[pos + 4:pos + 4 + ln]).decode()
        size, off, dt = struct.unpack_from("<QQi", mm, pos + 4 + ln)
        pos += 4 + ln + 20
        raw = mm[off:off + size]
        tensors[name] = raw.view(np.float32) if dt == F32 else raw
        dtypes[name] = dt
    return tensors, dtypes


def model_bytes(cfg):
    return sum(nbytes(r, c, d) for _, r, c, d in tensor_list(cfg))


def decode_weight_bytes(cfg):
    """Algorithmic HBM bytes of Linear weights read per decode step: every Linear incl. lm_head, not the embedding table."""
    return sum(nbytes(r, c, d) for n, r, c, d in tensor_list(cfg) if d == Q4_0 and n != Names(cfg["rope_type"]).embed)
