"""Seeded inputs for the per-op pin against the REFERENCE's own CPU Op classes (shared by make_golden3.py, which runs the
compiled reference, and tests/test_oracle_refops.py, which runs the C restatement and -- when built -- the reference live).

OpType numbers: mllm/OpDefined.hpp:10-134.  DataType: mllm/Types.hpp:63-97.  Shapes are (batch, head, sequence, dimension),
the argument order of Tensor::reshape; arrays are in BSHD memory order [b][s][h][d] (or [b][c][t][h][w] for BCTHW)."""
import numpy as np

OP = dict(SOFTMAX=3, ROPE=8, RMSNORM=11, LINEAR=14, EMBEDDING=17, LAYERNORM=26, CONVOLUTION2D=29, CONVOLUTION3D=30, MULTIMODALROPE=35,
          F_APPLY_VISIOROPE=109)
F32, Q4_0 = 0, 2
BSHD, BCTHW = 0, 3
VALID = 1


def cases(P):
    """-> dict case name -> dict(params, weights, inputs [(array, shape, dtype, ctype)], calls).  P = oracle port (for Q4_0 blocks)."""
    rng = np.random.default_rng(20260105)
    c = {}
    x = rng.standard_normal((1, 7, 1, 2048)).astype(np.float32)
    w = (1 + 0.02 * rng.standard_normal(2048)).astype(np.float32)
    c["rmsnorm"] = dict(params={"type": OP["RMSNORM"], "norm_size": 2048, "epsilon": 1e-6}, weights={".weight": (w, F32)},
                        inputs=[(x, (1, 1, 7, 2048), F32, BSHD)])
    c["rmsnorm_unit_offset"] = dict(params={"type": OP["RMSNORM"], "norm_size": 2048, "epsilon": 1e-5, "add_unit_offset": 1}, weights={".weight": ((w - 1).astype(np.float32), F32)},
                                    inputs=[(x, (1, 1, 7, 2048), F32, BSHD)])
    xh = rng.standard_normal((1, 5, 4, 128)).astype(np.float32)     # head-wise q/k norm (Qwen3): [B,S,H,D], norm over D
    wh = (1 + 0.05 * rng.standard_normal(128)).astype(np.float32)
    c["rmsnorm_headwise"] = dict(params={"type": OP["RMSNORM"], "norm_size": 128, "epsilon": 1e-6}, weights={".weight": (wh, F32)},
                                 inputs=[(xh, (1, 4, 5, 128), F32, BSHD)])
    xl = (rng.standard_normal((1, 9, 1, 1024)) * 3 + 0.5).astype(np.float32)
    wl = (1 + 0.1 * rng.standard_normal(1024)).astype(np.float32)
    bl = (0.1 * rng.standard_normal(1024)).astype(np.float32)
    c["layernorm"] = dict(params={"type": OP["LAYERNORM"], "norm_size": 1024, "bias": 1, "epsilon": 1e-5}, weights={".weight": (wl, F32), ".bias": (bl, F32)},
                          inputs=[(xl, (1, 1, 9, 1024), F32, BSHD)])
    c["layernorm_nobias"] = dict(params={"type": OP["LAYERNORM"], "norm_size": 1024, "bias": 0, "epsilon": 1e-6}, weights={".weight": (wl, F32)},
                                 inputs=[(xl, (1, 1, 9, 1024), F32, BSHD)])
    for rt, name in ((2, "rope_llama"), (4, "rope_hf")):
        xr = rng.standard_normal((2, 5, 3, 64)).astype(np.float32)
        # two calls: the op keeps h_cnt_ and the second call continues at position 5 (op/CPURoPE.cpp:517-530)
        c[name] = dict(params={"type": OP["ROPE"], "pose_type": rt, "rope_theta": 10000.0, "max_position_embeddings": 40}, weights=None,
                       inputs=[(xr, (2, 3, 5, 64), F32, BSHD)], calls=2)
    xr = rng.standard_normal((1, 6, 2, 128)).astype(np.float32)
    c["rope_hf_theta1e6_d128"] = dict(params={"type": OP["ROPE"], "pose_type": 4, "rope_theta": 1000000.0, "max_position_embeddings": 64}, weights=None,
                                      inputs=[(xr, (1, 2, 6, 128), F32, BSHD)], calls=2)
    xr = rng.standard_normal((1, 4, 2, 64)).astype(np.float32)
    c["rope_hf_partial"] = dict(params={"type": OP["ROPE"], "pose_type": 4, "rope_theta": 10000.0, "max_position_embeddings": 32, "partial_rotary_factor": 0.5},
                                weights=None, inputs=[(xr, (1, 2, 4, 64), F32, BSHD)])
    xr = rng.standard_normal((1, 6, 2, 128)).astype(np.float32)
    # LLaMA-3 frequency scaling (RoPEThetaType LLAMA3 = 1, op/CPURoPE.cpp:34-72); small original context so all three branches occur
    c["rope_hf_llama3"] = dict(params={"type": OP["ROPE"], "pose_type": 4, "rope_theta": 500000.0, "max_position_embeddings": 64, "rope_type": 1, "factor": 8.0,
                                       "low_freq_factor": 1.0, "high_freq_factor": 4.0, "original_max_position_embeddings": 64.0},
                               weights=None, inputs=[(xr, (1, 2, 6, 128), F32, BSHD)], calls=2)
    S, H, D = 11, 2, 128
    pos = np.stack([np.arange(S) + 3, 2 * np.arange(S), 3 * np.arange(S) + 1]).astype(np.float32)
    params = {"type": OP["MULTIMODALROPE"], "rope_theta": 1000000.0, "max_position_embeddings": 1024}
    for i, s_ in enumerate((16, 24, 24)):
        params[f"mrope_section_{i}"] = s_
    xm = rng.standard_normal((1, S, H, D)).astype(np.float32)
    c["mrope"] = dict(params=params, weights=None, inputs=[(xm, (1, H, S, D), F32, BSHD), (pos.reshape(3, 1, 1, S), (3, 1, 1, S), F32, BSHD)])
    S, H, D = 6, 2, 80
    xv = rng.standard_normal((1, S, H, D)).astype(np.float32)
    ang = rng.uniform(0, 60, (S, D // 2)).astype(np.float32)
    c["vision_rope"] = dict(params={"type": OP["F_APPLY_VISIOROPE"]}, weights=None,
                            inputs=[(xv, (1, H, S, D), F32, BSHD), (ang.reshape(1, S, 1, D // 2), (1, 1, S, D // 2), F32, BSHD)])
    tab = P.quantize_q4_0((rng.standard_normal((50, 256)) * 0.02).astype(np.float32))
    ids = rng.integers(0, 50, (2, 7)).astype(np.float32)
    ids[1, 3] = -1                                                   # negative id: row skipped (op/CPUEmbedding.cpp:44-46)
    c["embedding_q4_0"] = dict(params={"type": OP["EMBEDDING"], "hidden_size": 256, "vocab_size": 50}, weights={".weight": (tab, Q4_0)},
                               inputs=[(ids.reshape(2, 7, 1, 1), (2, 1, 7, 1), F32, BSHD)],
                               valid=(ids >= 0).reshape(2, 7, 1, 1))   # the skipped row is left as allocated (undefined) by the reference
    tabf = (rng.standard_normal((20, 64)) * 0.02).astype(np.float32)
    idf = rng.integers(0, 20, (1, 9)).astype(np.float32)
    c["embedding_f32"] = dict(params={"type": OP["EMBEDDING"], "hidden_size": 64, "vocab_size": 20}, weights={".weight": (tabf, F32)},
                              inputs=[(idf.reshape(1, 9, 1, 1), (1, 1, 9, 1), F32, BSHD)])
    xc = rng.standard_normal((1, 3, 28, 28)).astype(np.float32)      # mllm image tensor: head = height, sequence = channel, dimension = width
    wc = (rng.standard_normal((16, 3, 14, 14)) * 0.05).astype(np.float32)
    bc = (rng.standard_normal(16) * 0.1).astype(np.float32)
    c["conv2d_patch"] = dict(params={"type": OP["CONVOLUTION2D"], "kernal_h": 14, "kernal_w": 14, "stride_h": 14, "stride_w": 14, "in_channel": 3,
                                     "out_channel": 16, "padding": VALID, "bias": 1},
                             weights={".weight": (wc, F32), ".bias": (bc, F32)}, inputs=[(xc, (1, 28, 3, 28), F32, BSHD)])
    x3 = rng.standard_normal((1, 3, 2, 28, 28)).astype(np.float32)
    w3 = (rng.standard_normal((8, 3, 2, 14, 14)) * 0.05).astype(np.float32)
    c["conv3d_patch"] = dict(params={"type": OP["CONVOLUTION3D"], "kernal_t": 2, "kernal_h": 14, "kernal_w": 14, "stride_t": 2, "stride_h": 14,
                                     "stride_w": 14, "in_channel": 3, "out_channel": 8, "padding": VALID, "bias": 0},
                             weights={".weight": (w3, F32)}, inputs=[(x3, (1, 3, 2, 28, 28), F32, BCTHW)])
    sx = (rng.standard_normal((3, 7, 1, 9)) * 4).astype(np.float32)   # attention scores as the eager path hands them over: [B*H, 1, S, S']
    c["softmax"] = dict(params={"type": OP["SOFTMAX"], "axis": 3, "do_causal_mask": 0}, weights=None, inputs=[(sx, (3, 1, 7, 9), F32, BSHD)])
    Wl = P.quantize_q4_0((rng.standard_normal((96, 256)) * 0.02).astype(np.float32))
    bq = (rng.standard_normal(96) * 0.1).astype(np.float32)
    xq = rng.standard_normal((1, 5, 1, 256)).astype(np.float32)
    c["linear_q4_0_bias"] = dict(params={"type": OP["LINEAR"], "in_features": 256, "out_features": 96, "bias": 1}, weights={".weight": (Wl, Q4_0), ".bias": (bq, F32)},
                                 inputs=[(xq, (1, 1, 5, 256), F32, BSHD)])
    Wf = (rng.standard_normal((24, 64)) * 0.05).astype(np.float32)
    xf = rng.standard_normal((1, 3, 1, 64)).astype(np.float32)
    c["linear_f32"] = dict(params={"type": OP["LINEAR"], "in_features": 64, "out_features": 24, "bias": 0}, weights={".weight": (Wf, F32)},
                           inputs=[(xf, (1, 1, 3, 64), F32, BSHD)])
    return c


def run_reference(po, case):
    """Outputs of the compiled reference op: list (one per call) of fp32 arrays [b, s, h, d] (or [b, c, t, h, w])."""
    if case["params"]["type"] == OP["ROPE"]:
        # CPURoPE keeps its sin/cos tables in process-global statics and rebuilds them only when the rotated width grows or the
        # pose type changes (op/CPURoPE.cpp:13-18,207-224): one model per process.  Cases here run back to back in one process, so
        # flip the pose type with a throw-away op first; the case then always builds its own tables.
        other = 2 if case["params"]["pose_type"] == 4 else 4
        dummy = po.RefOp({"type": OP["ROPE"], "pose_type": other, "rope_theta": 10000.0, "max_position_embeddings": 4}, "dummy", None)
        dummy.run([(np.zeros((1, 1, 1, 8), dtype=np.float32), (1, 1, 1, 8), F32, BSHD)])
        dummy.close()
    op = po.RefOp(case["params"], "op", case["weights"])
    outs = [op.run(case["inputs"])[0] for _ in range(case.get("calls", 1))]
    op.close()
    return outs


def run_port(P, name, case):
    """The same cases through the C restatement (oracle.c)."""
    ins = case["inputs"]
    x = ins[0][0]
    pr = case["params"]
    w = case["weights"] or {}
    if name.startswith("rmsnorm"):
        return [P.only.rmsnorm(x.reshape(-1, x.shape[-1]), w[".weight"][0], pr["epsilon"], bool(pr.get("add_unit_offset", 0))).reshape(x.shape)]
    if name.startswith("layernorm"):
        return [P.only.layernorm(x.reshape(-1, x.shape[-1]), w[".weight"][0], w[".bias"][0] if ".bias" in w else None, pr["epsilon"]).reshape(x.shape)]
    if name == "rope_hf_llama3":
        return None  # the scaled-theta table has no C restatement: the CUDA op is checked against this golden directly (GPU tests)
    if name.startswith("rope_"):
        B, S, H, D = x.shape
        rot = int(D * pr.get("partial_rotary_factor", 1.0))
        s, c = P.only.rope_tables(pr["pose_type"], pr["max_position_embeddings"], rot, pr["rope_theta"])
        return [np.stack([P.only.rope_apply(pr["pose_type"], x[b], k * S, s, c, rot=rot) for b in range(B)]) for k in range(case.get("calls", 1))]
    if name == "mrope":
        pos = ins[1][0].reshape(3, -1)
        sec = [pr[f"mrope_section_{i}"] for i in range(3)]
        s, c = P.only.mrope_tables(pos, x.shape[-1], pr["rope_theta"], sec + sec)
        return [P.only.rope_apply(4, x[0], 0, s, c)[None]]
    if name == "vision_rope":
        return [P.only.rope_vision(x[0], ins[1][0].reshape(x.shape[1], -1))[None]]
    if name == "embedding_q4_0":
        B, S = x.shape[:2]
        return [P.only.embedding_q4_0(w[".weight"][0], x.reshape(B, S), pr["hidden_size"]).reshape(B, S, 1, -1)]
    if name == "embedding_f32":
        tab = w[".weight"][0]
        return [tab[x.reshape(-1).astype(np.int64)].reshape(x.shape[0], x.shape[1], 1, -1)]
    if name == "conv2d_patch":
        return [P.conv2d_valid(x, w[".weight"][0], w[".bias"][0], (pr["stride_h"], pr["stride_w"]))]
    if name == "conv3d_patch":
        return [P.only.conv3d_valid(x, w[".weight"][0], None, (pr["stride_t"], pr["stride_h"], pr["stride_w"]))]
    if name == "softmax":
        b, s, h, d = x.shape
        return [P.softmax(x.reshape(b * s * h, 1, d)).reshape(x.shape)]
    if name == "linear_q4_0_bias":
        return [P.linear(w[".weight"][0], x.reshape(-1, x.shape[-1]), bias=w[".bias"][0]).reshape(1, x.shape[1], 1, -1)]
    if name == "linear_f32":
        return None  # no restatement: fp32-weight Linear is vec_dot_fp32 per output (checked on the GPU side against the golden)
    raise KeyError(name)
