"""GPU parity tests of the tensor-core prefill/append attention (attention_tc.cu) against the CPU oracle
(flash_attention_2_forward restated, oracle/oracle.c) and against the CUDA-core kernel it replaces.
fp32 results: 1e-5 absolute on O(1) outputs (the reference's own op tests use 1e-3, test/cpu/CPUTest.hpp:116)."""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def dev(a):
    import torch
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


def with_tc(on, fn):
    old = os.environ.get("MLLM_CUDA_ATTN_TC")
    os.environ["MLLM_CUDA_ATTN_TC"] = "1" if on else "0"
    try:
        return fn()
    finally:
        if old is None:
            del os.environ["MLLM_CUDA_ATTN_TC"]
        else:
            os.environ["MLLM_CUDA_ATTN_TC"] = old


# ragged query counts, append (Sk > Sq, (Sk - Sq) % 4 == 0 as the reference requires), GQA, both head dims
@pytest.mark.parametrize("Sq,Sk,Hq,Hkv,D", [(16, 16, 2, 2, 64), (64, 64, 4, 4, 128), (70, 70, 4, 2, 64), (129, 129, 2, 1, 128),
                                              (48, 112, 4, 4, 128), (200, 264, 2, 2, 64), (17, 81, 8, 2, 128)])
def test_prefill_tc_vs_oracle(cu, oracle, Sq, Sk, Hq, Hkv, D):
    rng = np.random.default_rng(Sq * 7 + Sk)
    q = rng.standard_normal((1, Sq, Hq, D)).astype(np.float32)
    k = rng.standard_normal((1, Sk, Hkv, D)).astype(np.float16)
    v = rng.standard_normal((1, Sk, Hkv, D)).astype(np.float16)
    o = with_tc(True, lambda: cu.flash_attn_f16kv(dev(q), dev(k), dev(v)).cpu().numpy())
    want = oracle.flash_attn_f16kv(q[0], k[0].view(np.uint16), v[0].view(np.uint16))
    assert np.abs(o[0] - want).max() <= 1e-5 * max(1.0, np.abs(want).max())


def test_prefill_tc_sharp_softmax_and_large_q(cu, oracle):
    """Large |q| (scores of +-100s, near one-hot softmax) exercises the hi/lo split: an fp16-only q would miss 1e-5."""
    rng = np.random.default_rng(5)
    Sq = Sk = 96
    q = (rng.standard_normal((1, Sq, 4, 128)) * 12.0).astype(np.float32)
    k = rng.standard_normal((1, Sk, 4, 128)).astype(np.float16)
    v = (rng.standard_normal((1, Sk, 4, 128)) * 3.0).astype(np.float16)
    o = with_tc(True, lambda: cu.flash_attn_f16kv(dev(q), dev(k), dev(v)).cpu().numpy())
    want = oracle.flash_attn_f16kv(q[0], k[0].view(np.uint16), v[0].view(np.uint16))
    assert np.abs(o[0] - want).max() <= 2e-5 * max(1.0, np.abs(want).max())


def test_prefill_tc_7b_shape_matches_cuda_core_kernel(cu):
    """BASELINE configs[1] attention shape (S 512, 32 heads, D 128), batch 2 inside a cache with a larger limit."""
    import torch
    rng = np.random.default_rng(11)
    B, S, H, D, limit = 2, 512, 32, 128, 1024
    q = dev(rng.standard_normal((B, S, H, D)).astype(np.float32))
    kc = torch.zeros((B, limit, H, D), dtype=torch.float16, device="cuda")
    vc = torch.zeros_like(kc)
    kc[:, :S] = dev(rng.standard_normal((B, S, H, D)).astype(np.float16))
    vc[:, :S] = dev(rng.standard_normal((B, S, H, D)).astype(np.float16))
    run = lambda: cu.flash_attn_f16kv(q, kc[:, :S], vc[:, :S], kv_bstride=limit * H * D).cpu().numpy()
    a = with_tc(True, run)
    b = with_tc(False, run)
    assert np.isfinite(a).all()
    assert np.abs(a - b).max() <= 1e-5
    # size-independent property: the first query row attends to key 0 only -> o[0] == v[0] exactly widened
    assert np.array_equal(a[:, 0], vc[:, 0].float().cpu().numpy())


def test_noncausal_f16kv(cu):
    rng = np.random.default_rng(3)
    S, H, D = 150, 4, 64
    q = rng.standard_normal((1, S, H, D)).astype(np.float32)
    k = rng.standard_normal((1, S, H, D)).astype(np.float16)
    v = rng.standard_normal((1, S, H, D)).astype(np.float16)
    got = with_tc(True, lambda: cu.flash_attn_f16kv(dev(q), dev(k), dev(v), causal=False).cpu().numpy())[0]
    s = np.einsum("qhd,khd->hqk", q[0].astype(np.float64), k[0].astype(np.float64)) / np.sqrt(D)
    p = np.exp(s - s.max(-1, keepdims=True))
    p /= p.sum(-1, keepdims=True)
    want = np.einsum("hqk,khd->qhd", p, v[0].astype(np.float64))
    assert np.abs(got - want).max() <= 1e-5


# fp32 K/V (vision towers): CLIP ViT-L/14-336 (577 tokens, D 64), Qwen2-VL ViT (D 80), and D 128
@pytest.mark.parametrize("S,H,D", [(577, 16, 64), (333, 16, 80), (1369, 4, 80), (200, 4, 128)])
def test_f32kv_tc_vs_float64(cu, S, H, D):
    rng = np.random.default_rng(S + D)
    q, k, v = (rng.standard_normal((1, S, H, D)).astype(np.float32) for _ in range(3))
    got = with_tc(True, lambda: cu.flash_attn_f32kv(dev(q), dev(k), dev(v), causal=False).cpu().numpy())[0]
    old = with_tc(False, lambda: cu.flash_attn_f32kv(dev(q), dev(k), dev(v), causal=False).cpu().numpy())[0]
    s = np.einsum("qhd,khd->hqk", q[0].astype(np.float64), k[0].astype(np.float64)) / np.sqrt(D)
    p = np.exp(s - s.max(-1, keepdims=True))
    p /= p.sum(-1, keepdims=True)
    want = np.einsum("hqk,khd->qhd", p, v[0].astype(np.float64))
    assert np.abs(got - want).max() <= 1e-5
    assert np.abs(got - old).max() <= 1e-5
