"""GPU parity tests of the whole path through the host-side backend (CUDABackend, ops, ParamLoader,
LLaMA graph) with HOST buffers, against the CPU model oracle on the same random-init .mllm file:
per-layer hidden states, logits (max-abs and cosine) and greedy token ids over a fixed horizon."""
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


def make(tmp_path_factory, name, limit, seed):
    cfg = mf.config(name, cache_limit=limit)
    t = mf.random_tensors(cfg, seed=seed)
    path = str(tmp_path_factory.mktemp(name) / f"{name}.mllm")
    mf.write_model(path, cfg, t)
    return cfg, t, path


def cosine(a, b):
    a, b = a.astype(np.float64).ravel(), b.astype(np.float64).ravel()
    return float((a * b).sum() / (np.linalg.norm(a) * np.linalg.norm(b) + 1e-30))


@pytest.mark.parametrize("name,B,S", [("toy-llama", 1, 8), ("toy-gqa", 1, 12), ("toy-gqa", 2, 4), ("toy-llama", 1, 5)])
def test_prefill_then_decode_matches_oracle(be, oracle, tmp_path_factory, name, B, S):
    from oracle import pyoracle as po
    cfg, t, path = make(tmp_path_factory, name, 48, seed=S)
    rng = np.random.default_rng(B * 100 + S)
    ids = rng.integers(0, cfg["vocab"], (B, S)).astype(np.float32)
    cpu = po.ModelRunner(oracle, cfg, t, mf.Names(cfg["rope_type"]), max_batch=B)
    gpu = be.LlamaModel(cfg).load(path)
    lc, dc = cpu.forward(ids, dump=True)
    lg, dg = gpu.forward(ids, dump=True)
    # hidden state after the embedding is a pure dequantisation: bit-exact
    assert np.array_equal(dg[0].view(np.uint32), dc[0].view(np.uint32))
    for L in range(1, cfg["layers"] + 1):
        # fp32 sums are re-ordered (~1e-7 relative); a 1-ulp change can flip one Q8_0 activation rounding or
        # one fp16 KV rounding downstream (a ~1e-2 step on that element), so layers agree to ~1e-3 of max
        # Which roundings flip depends on the summation order of every kernel upstream, so the bar is split: the bulk of the
        # elements (99th percentile) within 2e-3 of max, the few that sit downstream of a flip within 1.5e-2.
        err, ref = np.abs(dg[L] - dc[L]), max(np.abs(dc[L]).max(), 1e-3)
        assert np.quantile(err, 0.99) <= 2e-3 * ref, f"layer {L}"
        assert err.max() <= 1.5e-2 * ref, f"layer {L}"
        assert cosine(dg[L], dc[L]) >= 0.99999, f"layer {L}"
    # logits: SURVEY 8(c) bar -- cosine >= 0.9999; max-abs <= 1.5e-2 * max|logit| (one flipped rounding, as above), bulk <= 2e-3
    assert cosine(lg, lc) >= 0.9999
    assert np.abs(lg - lc).max() <= 1.5e-2 * np.abs(lc).max()
    assert np.quantile(np.abs(lg - lc), 0.99) <= 2e-3 * np.abs(lc).max()
    for step in range(6):
        assert np.array_equal(lg.argmax(1), lc.argmax(1))
        nxt = lc.argmax(1).astype(np.float32)[:, None]
        lc, lg = cpu.forward(nxt), gpu.forward(nxt)
        assert cosine(lg, lc) >= 0.9999
    gpu.close()


def test_greedy_tokens_identical_over_horizon(be, oracle, tmp_path_factory):
    from oracle import pyoracle as po
    cfg, t, path = make(tmp_path_factory, "toy-gqa", 96, seed=21)
    rng = np.random.default_rng(4)
    ids = rng.integers(0, cfg["vocab"], (1, 16)).astype(np.float32)
    horizon = 64
    cpu = po.ModelRunner(oracle, cfg, t, mf.Names(cfg["rope_type"]))
    want, cur = [], ids
    for _ in range(horizon):
        tok = int(cpu.forward(cur).argmax())
        want.append(tok)
        cur = np.array([[tok]], dtype=np.float32)
    gpu = be.LlamaModel(cfg).load(path)
    got = gpu.generate(ids, horizon)[0].tolist()
    assert got == want
    # clear_kvcache + rerun reproduces the same tokens (KVCache::clearCache / RoPE::clearCache)
    gpu.clear_kvcache()
    assert gpu.generate(ids, 8)[0].tolist() == want[:8]
    gpu.close()


def test_cache_limit_is_an_error(be, tmp_path_factory):
    cfg, t, path = make(tmp_path_factory, "toy-llama", 8, seed=2)
    gpu = be.LlamaModel(cfg).load(path)
    ids = np.zeros((1, 12), dtype=np.float32)
    with pytest.raises(be.BackendError, match="cache limit"):
        gpu.forward(ids)
    gpu.close()


def test_unsupported_op_fails_loudly(be):
    with pytest.raises(be.BackendError, match="no CUDA implementation"):
        be.run_op({"type": 36.0}, "avgpool", [np.zeros((1, 1, 1, 4), dtype=np.float32)])


def test_per_op_rmsnorm_and_silu_through_backend(be, oracle, tmp_path_factory):
    """test/cpu/CPURMSNormTest.cpp / CPUSiluTest.cpp pattern: op built by opCreate, run by runOp."""
    rng = np.random.default_rng(7)
    x = rng.standard_normal((2, 1, 5, 256)).astype(np.float32)  # [b,h,s,d] logical == BSHD memory when h == 1
    w = (1 + 0.02 * rng.standard_normal(256)).astype(np.float32)
    p = str(tmp_path_factory.mktemp("op") / "norm.mllm")
    mf.write_mllm(p, {"n.weight": w}, {"n.weight": mf.F32})
    y = be.run_op({"type": 11.0, "norm_size": 256.0, "epsilon": 1e-6}, "n", [x], p)
    want = oracle.only.rmsnorm(x.reshape(10, 256), w, 1e-6)
    assert np.abs(y.reshape(10, 256).view(np.uint32).astype(np.int64) - want.view(np.uint32).astype(np.int64)).max() <= 1
    z = be.run_op({"type": 4.0}, "act", [x])
    assert np.array_equal(z.reshape(-1).view(np.uint32), oracle.silu(x.reshape(-1)).view(np.uint32))
