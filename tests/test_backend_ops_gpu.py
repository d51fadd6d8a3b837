"""GPU tests through the host-side backend (CUDABackend::opCreate + runOp, ParamLoader) for the rows added after the first
LLaMA slice: k-quant and Q4_0_4_4 model files end to end against the CPU model oracle, and the multimodal ops
(GELU / QuickGELU, SoftMax, Convolution2D, MultimodalRoPE, vision RoPE, transpose + mm) per op, the
test/cpu/CPUTest.hpp pattern."""
import numpy as np
import pytest

from mllm_b200 import mllm_file as mf

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def be():
    import torch
    assert torch.cuda.is_available()
    from mllm_b200 import backend, build
    build.build_all()
    backend.init_backend(0)
    return backend


def bits(a):
    return np.ascontiguousarray(a, dtype=np.float32).view(np.uint32)


def cosine(a, b):
    a, b = a.astype(np.float64).ravel(), b.astype(np.float64).ravel()
    return float((a * b).sum() / (np.linalg.norm(a) * np.linalg.norm(b) + 1e-30))


@pytest.mark.parametrize("name,wtype,B,S", [("toy-llama", mf.Q4_K, 1, 8), ("toy-gqa", mf.Q4_K, 2, 12), ("toy-llama", mf.Q4_0_4_4, 1, 9), ("toy-gqa", mf.Q4_0_4_4, 1, 16)])
def test_model_with_other_weight_formats(be, oracle, tmp_path_factory, name, wtype, B, S):
    """LLaVA's LLM is LLaMA with Q4_K Linears (config 3); the shipped checkpoints are Q4_0_4_4: the file is
    written in that dtype, the CPU oracle runs the reference arithmetic for it (Q4_0_4_4 == Q4_0 on de-interleaved rows)."""
    from oracle import pyoracle as po
    cfg = mf.config(name, cache_limit=48)
    cfg["wtype"] = wtype
    t = mf.random_tensors(cfg, seed=S + wtype)
    path = str(tmp_path_factory.mktemp(name) / f"{name}-{wtype}.mllm")
    mf.write_model(path, cfg, t)
    rng = np.random.default_rng(B * 100 + S)
    ids = rng.integers(0, cfg["vocab"], (B, S)).astype(np.float32)
    cpu = po.ModelRunner(oracle, cfg, t, mf.Names(cfg["rope_type"]), max_batch=B)
    gpu = be.LlamaModel(cfg).load(path)
    lc, dc = cpu.forward(ids, dump=True)
    lg, dg = gpu.forward(ids, dump=True)
    assert np.array_equal(bits(dg[0]), bits(dc[0]))
    for L in range(1, cfg["layers"] + 1):
        # as in test_model_gpu.py: re-ordered fp32 sums can flip one activation rounding downstream; a flipped Q8_K value moves an
        # output by d_x * d_w * scale (6-bit scale up to 63), hence the wider bar for k-quants (kernels themselves agree to 2e-5)
        # bulk of the elements (99th percentile) within 2e-3 of max; the few downstream of a flipped rounding within 1.5e-2
        err, ref = np.abs(dg[L] - dc[L]), max(np.abs(dc[L]).max(), 1e-3)
        assert np.quantile(err, 0.99) <= (4e-3 if wtype == mf.Q4_K else 2e-3) * ref, f"layer {L}"
        assert err.max() <= 1.5e-2 * ref, f"layer {L}"
        assert cosine(dg[L], dc[L]) >= 0.99999, f"layer {L}"
    assert cosine(lg, lc) >= 0.9999
    for step in range(6):
        # random-init toy models have near-flat logits: the greedy pick must agree whenever the CPU's own top-2 margin is not
        # inside the rounding-flip noise (a flipped Q8_K activation, 1/127 of a 256-wide block's max, cascades to ~1e-2 relative)
        top2 = np.sort(lc, axis=1)[:, -2:]
        clear = (top2[:, 1] - top2[:, 0]) > 2e-2 * np.abs(lc).max()
        assert np.array_equal(lg.argmax(1)[clear], lc.argmax(1)[clear])
        nxt = lc.argmax(1).astype(np.float32)[:, None]
        lc, lg = cpu.forward(nxt), gpu.forward(nxt)
        # single-token steps replay the captured decode graph for every weight format (k-quants: op-level kernels on fixed buffers)
        assert gpu.last_step_fused()
        assert cosine(lg, lc) >= (0.9995 if wtype == mf.Q4_K else 0.9999)
    gpu.close()


@pytest.mark.parametrize("optype,kind", [(24.0, 0), (25.0, 1)])
def test_gelu_ops_bit_exact(be, oracle, optype, kind):
    rng = np.random.default_rng(kind)
    x = (rng.standard_normal((1, 1, 77, 1024)) * 3).astype(np.float32)
    y = be.run_op({"type": optype}, "act", [x])
    want = oracle.only.gelu(oracle.gelu_table(kind), x.reshape(-1))
    assert np.array_equal(bits(y.reshape(-1)), bits(want))


def test_softmax_op_causal(be, oracle):
    rng = np.random.default_rng(3)
    x = (rng.standard_normal((8, 1, 16, 24)) * 3).astype(np.float32)  # [G,1,S,D]: 8 past keys + 16 new
    y = be.run_op({"type": 3.0, "axis": 3.0, "do_causal_mask": 1.0}, "sm", [x])
    want = oracle.softmax(x.reshape(8, 16, 24), True, 8)
    assert np.abs(bits(y.reshape(8, 16, 24)).astype(np.int64) - bits(want).astype(np.int64)).max() <= 4


def test_convolution2d_op(be, oracle, tmp_path_factory):
    rng = np.random.default_rng(4)
    x = rng.standard_normal((1, 3, 56, 42)).astype(np.float32)        # NCHW
    w = (rng.standard_normal((32, 3, 14, 14)) * 0.05).astype(np.float32)
    b = rng.standard_normal(32).astype(np.float32)
    p = str(tmp_path_factory.mktemp("op") / "conv.mllm")
    mf.write_mllm(p, {"pe.weight": w, "pe.bias": b}, {"pe.weight": mf.F32, "pe.bias": mf.F32})
    # mllm shape [b, h=height, s=channel, d=width]; BSHD memory of that shape is NCHW
    y = be.run_op({"type": 29.0, "in_channel": 3.0, "out_channel": 32.0, "kernal_h": 14.0, "kernal_w": 14.0, "stride_h": 14.0, "stride_w": 14.0,
                   "padding": 1.0, "bias": 1.0}, "pe", [x.reshape(1, 3, 56, 42)], p, shapes=[(1, 56, 3, 42)])
    want = oracle.conv2d_valid(x, w, b, (14, 14))
    assert y.shape == (1, 32, 4, 3)  # BSHD memory order: [b, s=O, h=oh, d=ow]
    assert np.abs(y - want).max() <= 2e-5 * np.abs(want).max()


def test_multimodal_rope_op_bit_exact(be, oracle):
    rng = np.random.default_rng(5)
    S, H, D = 19, 4, 128
    x = rng.standard_normal((1, S, H, D)).astype(np.float32)  # BSHD memory
    pos = np.stack([np.arange(S) // 2, rng.integers(0, 30, S), rng.integers(0, 30, S)]).astype(np.float32)
    y = be.run_op({"type": 35.0, "rope_theta": 1e6, "max_position_embeddings": 32768.0, "mrope_section_0": 16.0, "mrope_section_1": 24.0,
                   "mrope_section_2": 24.0}, "q_rope", [x, pos.reshape(3, 1, 1, S)], shapes=[(1, H, S, D), (3, 1, 1, S)])
    s, c = oracle.only.mrope_tables(pos, D, 1e6, [16, 24, 24])
    want = oracle.only.rope_apply(4, x[0], 0, s, c)
    assert np.array_equal(bits(y[0]), bits(want))


def test_vision_rope_op(be, oracle):
    rng = np.random.default_rng(6)
    S, H, D = 64, 16, 80
    x = rng.standard_normal((1, S, H, D)).astype(np.float32)
    ang = rng.uniform(0, 70, (S, D // 2)).astype(np.float32)
    y = be.run_op({"type": 109.0}, "vrope", [x, ang.reshape(1, S, 1, D // 2)], shapes=[(1, H, S, D), (1, 1, S, D // 2)])
    assert np.abs(y[0] - oracle.only.rope_vision(x[0], ang)).max() <= 2e-6 * np.abs(x).max()


def test_transpose_then_mm_is_qk(be):
    """eager spelling: q.transpose(HEAD, SEQUENCE) and k likewise, then mm(q, k^T) -> [B*H,1,Sq,Sk] scores"""
    rng = np.random.default_rng(7)
    S, H, D = 10, 4, 32
    q = rng.standard_normal((1, S, H, D)).astype(np.float32)
    k = rng.standard_normal((1, S, H, D)).astype(np.float32)
    qt = be.run_op({"type": 92.0}, "qt", [q], shapes=[(1, H, S, D)])
    kt = be.run_op({"type": 92.0}, "kt", [k], shapes=[(1, H, S, D)])
    assert np.array_equal(qt.reshape(H, S, D), q[0].transpose(1, 0, 2))
    sc = be.run_op({"type": 87.0, "transpose0": 0.0, "transpose1": 1.0}, "qk", [qt.reshape(H, S, 1, D), kt.reshape(H, S, 1, D)],
                   shapes=[(H, 1, S, D), (H, 1, S, D)])
    want = np.einsum("hqd,hkd->hqk", q[0].transpose(1, 0, 2).astype(np.float64), k[0].transpose(1, 0, 2).astype(np.float64))
    assert np.abs(sc.reshape(H, S, S) - want).max() <= 1e-5 * np.abs(want).max()


def test_parameter_op_is_the_loaded_tensor(be, tmp_path_factory):
    """ViT cls_token / position_embeddings (models/vit/modeling_vit.hpp:66-84): Parameter(batch, seq, head, dim, name)."""
    rng = np.random.default_rng(8)
    w = rng.standard_normal((1, 5, 1, 64)).astype(np.float32)   # BSHD memory
    p = str(tmp_path_factory.mktemp("op") / "param.mllm")
    mf.write_mllm(p, {"embeddings.position_embedding": w}, {"embeddings.position_embedding": mf.F32})
    y = be.run_op({"type": 1.0, "batch": 1.0, "seq": 5.0, "head": 1.0, "dim": 64.0}, "embeddings.position_embedding", [], p)
    assert np.array_equal(bits(y), bits(w))


@pytest.mark.parametrize("axis", [0, 1, 2, 3])
def test_cat_op_every_axis(be, axis):
    """Tensor::cat along BATCH / SEQUENCE / HEAD / DIMENSION (Chl 0/1/2/3): cls token + patches is the SEQUENCE case."""
    rng = np.random.default_rng(9 + axis)
    shp = [2, 7, 3, 16]  # BSHD memory order (b, s, h, d) == Chl order
    parts = []
    for n in (1, 4, 2):
        s = list(shp)
        s[axis] = n
        parts.append(rng.standard_normal(s).astype(np.float32))
    logical = [(a.shape[0], a.shape[2], a.shape[1], a.shape[3]) for a in parts]  # (b, h, s, d)
    y = be.run_op({"type": 90.0, "axis": float(axis)}, "cat", parts, shapes=logical)
    assert np.array_equal(bits(y), bits(np.concatenate(parts, axis=axis)))


def test_clip_vit_block_through_backend_ops(be, oracle, tmp_path_factory):
    """BASELINE configs[2] vision tower, one ViTBlock (models/vit/modeling_vit.hpp:31-62 with CLIP's QuickGELU and biased
    LayerNorm / Linear): norm1 -> q,k,v Linear (Q4_K weights, as the LLaVA file quantises them) -> non-causal flash attention
    over fp32 K/V (no cache) -> o Linear -> +x -> norm2 -> up Linear -> QuickGELU -> down Linear -> +tmp, every op created
    through CUDABackend::opCreate / runOp, against the same composition of the CPU oracle's functions."""
    from oracle import pyoracle as po
    rng = np.random.default_rng(21)
    S, Hd, H, F = 77, 256, 4, 512
    D = Hd // H
    x = rng.standard_normal((1, S, 1, Hd)).astype(np.float32)
    t, dt = {}, {}
    for n, (rows, cols) in {"q_proj": (Hd, Hd), "k_proj": (Hd, Hd), "v_proj": (Hd, Hd), "out_proj": (Hd, Hd), "fc1": (F, Hd), "fc2": (Hd, F)}.items():
        t[n + ".weight"] = mf.quantize_q4_K_np((0.05 * rng.standard_normal((rows, cols))).astype(np.float32))
        t[n + ".bias"] = (0.02 * rng.standard_normal(rows)).astype(np.float32)
        dt[n + ".weight"], dt[n + ".bias"] = mf.Q4_K, mf.F32
    for n in ("layer_norm1", "layer_norm2"):
        t[n + ".weight"] = (1 + 0.02 * rng.standard_normal(Hd)).astype(np.float32)
        t[n + ".bias"] = (0.02 * rng.standard_normal(Hd)).astype(np.float32)
        dt[n + ".weight"] = dt[n + ".bias"] = mf.F32
    path = str(tmp_path_factory.mktemp("vit") / "block.mllm")
    mf.write_mllm(path, t, dt)

    row = [(1, 1, S, Hd)]  # logical [b, h, s, d] of an activation whose BSHD memory is [S][hidden]

    def lin(name, a, n_in, n_out):
        return be.run_op({"type": 14.0, "in_features": float(n_in), "out_features": float(n_out), "bias": 1.0}, name, [a], path, shapes=[(1, 1, S, n_in)])

    def ln(name, a):
        return be.run_op({"type": 26.0, "norm_size": float(Hd), "bias": 1.0, "epsilon": 1e-5}, name, [a], path, shapes=row)

    # ---- GPU: op by op through the backend
    h1 = ln("layer_norm1", x)
    q, k, v = (lin(n, h1, Hd, Hd) for n in ("q_proj", "k_proj", "v_proj"))
    hs = [(1, H, S, D)] * 3  # view(-1, heads, -1, head_dim): same bytes, BSHD memory [S][H][D]
    a = be.run_op({"type": 110.0, "causal_mask": 0.0}, "fa2", [q, k, v], shapes=hs)
    o = lin("out_proj", a, Hd, Hd)
    tmp = be.run_op({"type": 83.0}, "add1", [o, x], shapes=row * 2)
    h2 = ln("layer_norm2", tmp)
    u = be.run_op({"type": 25.0}, "act", [lin("fc1", h2, Hd, F)], shapes=[(1, 1, S, F)])
    y = be.run_op({"type": 83.0}, "add2", [lin("fc2", u, F, Hd), tmp], shapes=row * 2)

    # ---- CPU oracle: the same composition
    def olin(name, a2):
        return oracle.linear_kq(t[name + ".weight"], a2, t[name + ".bias"], po.Q4_K)

    x2 = x.reshape(S, Hd)
    c1 = oracle.only.layernorm(x2, t["layer_norm1.weight"], t["layer_norm1.bias"], 1e-5)
    cq, ck, cv = (olin(n, c1).reshape(S, H, D) for n in ("q_proj", "k_proj", "v_proj"))
    sc = np.einsum("qhd,khd->hqk", cq.astype(np.float64), ck.astype(np.float64)) / np.sqrt(D)
    p = np.exp(sc - sc.max(-1, keepdims=True))
    p /= p.sum(-1, keepdims=True)
    ca = np.einsum("hqk,khd->qhd", p, cv.astype(np.float64)).astype(np.float32).reshape(S, Hd)
    ctmp = olin("out_proj", ca) + x2
    c2 = oracle.only.layernorm(ctmp, t["layer_norm2.weight"], t["layer_norm2.bias"], 1e-5)
    cu_ = oracle.only.gelu(oracle.gelu_table(1), olin("fc1", c2).reshape(-1)).reshape(S, F)
    want = olin("fc2", cu_) + ctmp

    # stage by stage first (each stage fed by the GPU's own previous stage would hide nothing; here both chains run end to end)
    assert np.abs(h1.reshape(S, Hd) - c1).max() <= 1e-5 * np.abs(c1).max()
    assert np.abs(q.reshape(S, H, D) - cq).max() <= 2e-5 * np.abs(cq).max()
    err, ref = np.abs(y.reshape(S, Hd) - want), np.abs(want).max()
    assert np.quantile(err, 0.99) <= 2e-3 * ref and err.max() <= 1.5e-2 * ref   # a flipped Q8_K rounding downstream, as in the model tests
    assert cosine(y.reshape(-1), want.reshape(-1)) >= 0.99999


@pytest.mark.parametrize("axis", [1, 2, 3])
def test_split_op(be, axis):
    """Tensor::split(each_dims, SEQUENCE / HEAD / DIMENSION): each piece is the matching slice (CPUsplitFunction)."""
    rng = np.random.default_rng(30 + axis)
    x = rng.standard_normal((2, 9, 4, 24)).astype(np.float32)  # BSHD memory order == Chl order (b, s, h, d)
    n = x.shape[axis]
    each = [1, n - 3, 2]
    params = {"type": 100.0, "num_splits": 3.0, "dim_0": float(each[0]), "dim_1": float(each[1]), "dim_2": float(each[2]), "split_dim": float(axis),
              "head_size": 1.0, "__n_out": 3.0}
    at = 0
    for i, e in enumerate(each):
        y = be.run_op(dict(params, __out_index=float(i)), "split", [x], shapes=[(2, 4, 9, 24)])
        want = np.take(x, range(at, at + e), axis=axis)
        assert y.shape == want.shape and np.array_equal(bits(y), bits(want))
        at += e


def test_convert_fp_data_half_and_back(be):
    """Backend::convert_fp_data (Tensor::half() / fp32()): fp32 -> fp16 rounds to nearest even like the reference's F16C conversion,
    fp16 -> fp32 is exact."""
    rng = np.random.default_rng(44)
    x = (rng.standard_normal((7, 96)) * np.float32(10.0) ** rng.integers(-6, 5, (7, 96))).astype(np.float32)
    x[0, :4] = [0.0, -0.0, 65504.0, 1e-8]
    x[1, 0] = np.float32(1.0 + 2.0 ** -11)  # a tie: rounds to even
    h, b = be.convert_fp_roundtrip(x)
    want = x.astype(np.float16)
    assert np.array_equal(h, want.view(np.uint16))
    assert np.array_equal(bits(b), bits(want.astype(np.float32)))
