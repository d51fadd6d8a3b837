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
Q4_K, Q6_K, Q4_0_4_4 = 12, 14, 19
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
    # rms_eps: the reference's TinyLLaMABlock hard-codes RMSNorm(hidden_dim, 1e-6, ...) (models/tinyllama/modeling_tinyllama.hpp:31-32)
    "tinyllama-1.1b": dict(vocab=32000, hidden=2048, layers=22, heads=32, kv_heads=4, ffn=5632, rope_type=ROPE_HF, rope_theta=10000.0, rms_eps=1e-6, qkv_bias=0),
    "llama2-7b": dict(vocab=32000, hidden=4096, layers=32, heads=32, kv_heads=32, ffn=11008, rope_type=ROPE_LLAMA, rope_theta=10000.0, rms_eps=1e-6, qkv_bias=0),
    # the LLM of Qwen2-VL-2B (models/qwen/configuration_qwen.hpp:167-181); text-only positions here (M-RoPE reduces to HF RoPE when t = h = w)
    "qwen2-vl-2b-llm": dict(vocab=151936, hidden=1536, layers=28, heads=12, kv_heads=2, ffn=8960, rope_type=ROPE_HF, rope_theta=1000000.0, rms_eps=1e-6, qkv_bias=1),
    "qwen1.5-1.8b": dict(vocab=151936, hidden=2048, layers=24, heads=16, kv_heads=16, ffn=5504, rope_type=ROPE_HF, rope_theta=1000000.0, rms_eps=1e-6, qkv_bias=1),
    # small graphs with the same structure, for parity tests the CPU oracle finishes in seconds
    "toy-llama": dict(vocab=512, hidden=256, layers=2, heads=4, kv_heads=4, ffn=512, rope_type=ROPE_LLAMA, rope_theta=10000.0, rms_eps=1e-6, qkv_bias=0),
    "toy-d128": dict(vocab=320, hidden=256, layers=2, heads=2, kv_heads=1, ffn=512, rope_type=ROPE_LLAMA, rope_theta=10000.0, rms_eps=1e-6, qkv_bias=0),
    # ffn = 128 x odd: Q4_0 rows of 18 * ffn / 32 bytes are 8 mod 16, as in Qwen1.5-1.8B (ffn 5504)
    "toy-qwen": dict(vocab=320, hidden=256, layers=2, heads=2, kv_heads=2, ffn=384, rope_type=ROPE_HF, rope_theta=1000000.0, rms_eps=1e-6, qkv_bias=1),
    # hidden = 128 x odd: Q4_0 rows over K = hidden are 8 mod 16 bytes, which the fused prologue GEMV refuses for 3+ rows (rc -2):
    # the decode plan must fall back op by op without losing the on-device pick
    "toy-h384": dict(vocab=320, hidden=384, layers=2, heads=3, kv_heads=3, ffn=512, rope_type=ROPE_HF, rope_theta=10000.0, rms_eps=1e-6, qkv_bias=0),
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
    """[(name, rows, cols, dtype)] of every parameter, in file order.  cfg["wtype"] (default Q4_0) is the Linear dtype --
    Q4_K for the LLaVA config, Q4_0_4_4 for the shipped interleaved files; the embedding table stays Q4_0 as the reference
    quantizer forces (tools/quantizer/QuantWriter.cpp:32-35,133)."""
    n = Names(cfg["rope_type"])
    H, D = cfg["hidden"], cfg["head_dim"]
    out = _tensor_list_q4_0(cfg, n, H, D)
    wt = cfg.get("wtype", Q4_0)
    if wt != Q4_0:
        out = [(nm, r, c, (wt if (d == Q4_0 and nm != n.embed) else d)) for nm, r, c, d in out]
    return out


def _tensor_list_q4_0(cfg, n, H, D):
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
    if dtype == F32:
        return rows * cols * 4
    if dtype == Q4_K:
        return rows * (cols // 256) * 144
    if dtype == Q6_K:
        return rows * (cols // 256) * 210
    return rows * (cols // 32) * 18


def quantize_q4_K_np(w):
    """A simple min/max Q4_K quantiser (NOT the reference's make_qkx2 search): produces valid block_q4_K bytes
    (mllm/DataType.hpp:93-98) for random-init model files; parity never depends on how the blocks were chosen, only on how
    they are consumed."""
    w = np.ascontiguousarray(w, dtype=np.float32)
    rows, cols = w.shape
    b = w.reshape(rows, cols // 256, 8, 32)
    mn = np.minimum(b.min(axis=3), 0)
    mx = b.max(axis=3)
    dsb = (mx - mn) / 15.0
    msb = -mn
    d = (dsb.max(axis=2) / 63.0).astype(np.float16)
    dm = (msb.max(axis=2) / 63.0).astype(np.float16)
    df, dmf = d.astype(np.float32), dm.astype(np.float32)
    with np.errstate(divide="ignore", invalid="ignore"):
        sc = np.where(df[..., None] > 0, np.rint(dsb / df[..., None]), 0).clip(0, 63).astype(np.uint8)
        m6 = np.where(dmf[..., None] > 0, np.rint(msb / dmf[..., None]), 0).clip(0, 63).astype(np.uint8)
        eff = df[..., None] * sc
        q = np.where(eff[..., None] > 0, np.rint((b + (dmf[..., None] * m6)[..., None]) / eff[..., None]), 0).clip(0, 15).astype(np.uint8)
    out = np.zeros((rows, cols // 256, 144), dtype=np.uint8)
    out[..., 0:2] = d.view(np.uint8).reshape(rows, cols // 256, 2)
    out[..., 2:4] = dm.view(np.uint8).reshape(rows, cols // 256, 2)
    # 6-bit packing, inverse of get_scale_min_k4 (ggml/QuantizeQ4.cpp:177-184)
    for j in range(4):
        out[..., 4 + j] = (sc[..., j] & 63) | ((sc[..., j + 4] >> 4) << 6)
        out[..., 8 + j] = (m6[..., j] & 63) | ((m6[..., j + 4] >> 4) << 6)
        out[..., 12 + j] = (sc[..., j + 4] & 0xF) | ((m6[..., j + 4] & 0xF) << 4)
    for j in range(4):
        out[..., 16 + 32 * j:16 + 32 * (j + 1)] = q[..., 2 * j, :] | (q[..., 2 * j + 1, :] << 4)
    return out.reshape(rows, cols // 256 * 144)


def pack_q4_0_4x4_np(rows_q4_0, cols):
    """plain block_q4_0 rows -> block_q4_0x4 groups (make_block_q4_0x4, interleave 4, XOR 0x88; ggml/GemmPack.cpp:45-60, :308-344)"""
    r = np.ascontiguousarray(rows_q4_0, dtype=np.uint8)
    n, nb = r.shape[0], cols // 32
    b = r.reshape(n // 4, 4, nb, 18)
    out = np.empty((n // 4, nb, 72), dtype=np.uint8)
    out[..., 0:8] = b[..., 0:2].transpose(0, 2, 1, 3).reshape(n // 4, nb, 8)
    qs = b[..., 2:].reshape(n // 4, 4, nb, 4, 4)  # [g, row, x, chunk, byte]
    out[..., 8:] = (qs.transpose(0, 2, 3, 1, 4) ^ 0x88).reshape(n // 4, nb, 64)
    return out.reshape(n, nb * 18)


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
        elif dt == Q4_K:
            t[name] = quantize_q4_K_np((0.02 * rng.standard_normal((rows, cols))).astype(np.float32))
        else:  # Q4_0, and Q4_0_4_4 (kept as plain rows here; write_model interleaves them for the file)
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
    tl = tensor_list(cfg)
    dt = {name: d for name, _, _, d in tl}
    cols = {name: c for name, _, c, _ in tl}
    return write_mllm(path, {n: (pack_q4_0_4x4_np(tensors[n], cols[n]) if dt[n] == Q4_0_4_4 else tensors[n]) for n in dt}, dt)


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
    return sum(nbytes(r, c, d) for n, r, c, d in tensor_list(cfg) if d != F32 and n != Names(cfg["rope_type"]).embed)


def write_random_model_torch(cfg, path, device="cuda", seed=1234):
    """Random-init .mllm of a BASELINE-sized model, generated with torch on `device` (plumbing: the 7B needs 6.6 G normals)
    and quantised with quantize_row_q4_0_reference's arithmetic (ggml/QuantizeQ4.cpp:33-68), written in the reference's container
    format.  W ~ N(0, 0.02^2) -> Q4_0, norm weights 1 + N(0, 0.02^2), biases N(0, 0.02^2).  Reused when the file
    already has the expected size."""
    import torch
    want = 12 + sum(4 + len(n.encode()) + 20 for n, *_ in tensor_list(cfg)) + model_bytes(cfg)
    if os.path.exists(path) and os.path.getsize(path) == want:
        return path
    g = torch.Generator(device=device).manual_seed(seed)

    def q4_0(rows, cols):
        out = np.empty((rows, cols // 32 * 18), dtype=np.uint8)
        step = max(1, (1 << 26) // cols)
        for r0 in range(0, rows, step):
            r = min(step, rows - r0)
            w = torch.randn((r, cols // 32, 32), generator=g, device=device) * 0.02
            idx = w.abs().argmax(dim=2, keepdim=True)
            mx = torch.gather(w, 2, idx)
            d = mx / -8.0
            idv = torch.where(d != 0, 1.0 / d, torch.zeros_like(d))
            q = torch.clamp(((w * idv) + 8.5).to(torch.int8), max=15).to(torch.uint8)
            blk = torch.empty((r, cols // 32, 18), dtype=torch.uint8, device=device)
            blk[..., 0:2] = d.to(torch.float16).view(torch.uint8).reshape(r, cols // 32, 2)
            blk[..., 2:] = q[..., :16] | (q[..., 16:] << 4)
            out[r0:r0 + r] = blk.reshape(r, -1).cpu().numpy()
        return out

    tensors, dts = {}, {}
    for name, rows, cols, dt in tensor_list(cfg):
        dts[name] = dt
        if dt == F32:
            base = 1.0 if "norm" in name else 0.0
            tensors[name] = (base + 0.02 * torch.randn(cols, generator=g, device=device)).cpu().numpy().astype(np.float32)
        elif dt == Q4_0:
            tensors[name] = q4_0(rows, cols)
        else:
            raise ValueError("write_random_model_torch: Q4_0 model files only")
    tmp = path + ".tmp"
    write_mllm(tmp, tensors, dts)
    os.replace(tmp, path)
    return path
