"""GPU parity tests for SURVEY 8a rows a4, a5, a7, a10, a11, a15, a17, a18: k-quant Linear, the Q4_0_4_4 load-time
de-interleave, eager softmax, M-RoPE / vision RoPE, GELU tables, patch-embed convolutions, fp32-KV (ViT) attention and the
strided copy -- every C-ABI kernel against the CPU oracle on the same seeded inputs and against the committed golden
vectors (tests/golden/ref_vectors2.npz, generated from the compiled reference).  Bit-exact for bytes / integers / table
lookups; stated tolerances for fp32 accumulations."""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def bits(a):
    return np.ascontiguousarray(a, dtype=np.float32).view(np.uint32)


def dev(a):
    import torch
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


@pytest.fixture(scope="module")
def g2():
    return np.load(os.path.join(ROOT, "tests", "golden", "ref_vectors2.npz"))


def rand_q4k(rng, N, K):
    W = rng.integers(0, 256, (N, K // 256, 144), dtype=np.uint8)
    W[:, :, 0:2] = (rng.uniform(0.001, 0.02, (N, K // 256, 1)).astype(np.float16)).view(np.uint8).reshape(N, K // 256, 2)
    W[:, :, 2:4] = (rng.uniform(0.001, 0.01, (N, K // 256, 1)).astype(np.float16)).view(np.uint8).reshape(N, K // 256, 2)
    return W.reshape(N, -1)


def rand_q6k(rng, N, K):
    W = rng.integers(0, 256, (N, K // 256, 210), dtype=np.uint8)
    W[:, :, 208:210] = (rng.uniform(0.0005, 0.005, (N, K // 256, 1)).astype(np.float16)).view(np.uint8).reshape(N, K // 256, 2)
    return W.reshape(N, -1)


def act_rows(rng, rows, k):
    x = (rng.standard_normal((rows, k)) * rng.uniform(0.01, 8, (rows, 1))).astype(np.float32)
    if rows > 2:
        x[1, :256] = 0.0
        x[2, :256] = np.arange(256, dtype=np.float32) * 0.25 - 32.0
        x[2, 5], x[2, 9] = 40.5, -40.5
    return x


# ------------------------------------------------------------------ a5: k-quants
def q8k_equal(got, want):
    got = got.reshape(-1, 292)
    want = want.reshape(-1, 292)
    zero = np.all(want[:, :260] == 0, axis=1)  # all-zero super-block: the reference leaves bsums unwritten
    return np.array_equal(got[~zero], want[~zero]) and np.array_equal(got[zero][:, :260], want[zero][:, :260])


def test_q8_K_blocks_bit_exact_golden(cu, g2):
    assert q8k_equal(cu.quantize_q8_K(dev(g2["q8k_x"])).cpu().numpy(), g2["q8k_blocks"])


@pytest.mark.parametrize("rows,k", [(1, 256), (1, 4096), (5, 11008), (64, 2048)])
def test_q8_K_blocks_bit_exact(cu, oracle, rows, k):
    rng = np.random.default_rng(rows * 7 + k)
    x = act_rows(rng, rows, k)
    assert q8k_equal(cu.quantize_q8_K(dev(x)).cpu().numpy(), oracle.quantize_q8_K(x))


def test_q4_K_integer_super_block_sums_bit_exact(cu, oracle):
    rng = np.random.default_rng(3)
    M, N, K = 3, 17, 1024
    W = rand_q4k(rng, N, K)
    xq = oracle.quantize_q8_K(act_rows(rng, M, K))
    a, b = cu.block_isums_q4_K_q8_K(dev(W), dev(xq), K)
    a, b = a.cpu().numpy(), b.cpu().numpy()
    for n in range(N):
        for m in range(M):
            wa, wb = oracle.only.block_isums_q4_K_q8_K(W[n], xq[m], K)
            assert np.array_equal(a[n, m], wa) and np.array_equal(b[n, m], wb)


def kq_check(y, want):
    # same products, different fp32 order than the 8 AVX lanes: <= 1e-3 * max|y| is the bar; measured ~1e-6
    y, want = y.astype(np.float64), want.astype(np.float64)
    assert np.abs(y - want).max() <= 2e-5 * max(np.abs(want).max(), 1e-6)
    assert (y * want).sum() / (np.linalg.norm(y) * np.linalg.norm(want) + 1e-30) >= 0.99999


def test_linear_q4_K_q6_K_golden(cu, g2):
    y = cu.linear(dev(g2["q4k_blocks"]), dev(g2["q8k_x"]), dev(g2["kq_bias"]), wtype=cu.TYPE_Q4_K).cpu().numpy()
    kq_check(y, g2["lin_q4k"])
    y = cu.linear(dev(g2["q6k_blocks"]), dev(g2["q8k_x"]), wtype=cu.TYPE_Q6_K).cpu().numpy()
    kq_check(y, g2["lin_q6k"])


# rows >= 9 with N % 16 == 0 take the tensor-core kernel (gemm_mma_q4k.cu: u8 x s8 mma.sync, integer scale / min sums): short K (4 K-splits x
# 128 rows), K = 2048 / 4096 (one stage per tile), K = 11008 (three stages per tile, the last one short), several row chunks, a partial chunk
@pytest.mark.parametrize("M,N,K", [(1, 64, 256), (1, 4096, 4096), (3, 50, 11008), (8, 33, 1024), (9, 130, 512), (70, 200, 2048), (130, 64, 4096),
                                   (9, 128, 512), (70, 208, 2048), (200, 96, 11008), (64, 2048, 4096), (608, 64, 1024), (577, 48, 4096)])
@pytest.mark.parametrize("wt", ["q4k", "q6k"])
def test_linear_k_quants(cu, oracle, M, N, K, wt):
    rng = np.random.default_rng(M * 5 + N + K)
    W = rand_q4k(rng, N, K) if wt == "q4k" else rand_q6k(rng, N, K)
    wtype = cu.TYPE_Q4_K if wt == "q4k" else cu.TYPE_Q6_K
    x = act_rows(rng, M, K)
    b = (rng.standard_normal(N) * 0.02).astype(np.float32)
    y = cu.linear(dev(W), dev(x), dev(b), wtype=wtype).cpu().numpy()
    kq_check(y, oracle.linear_kq(W, x, b, wtype))


# ---- the single-token chain of a Q4_K body: ops folded into the activation quantiser, matrices sharing one input in one launch, the
# residual in the GEMV epilogue.  The quantised bytes must be those of the op-by-op sequence (held against the oracle above); the GEMV keeps
# the exact integers per lane and differs from the op-level Linear only in fp32 summation order (kq_check's bar).
@pytest.mark.parametrize("rows,k", [(1, 256), (1, 4096), (3, 11008), (8, 2048)])
def test_prep_q8_K_equals_separate_ops(cu, rows, k):
    import torch
    rng = np.random.default_rng(rows + k)
    x, u = dev(act_rows(rng, rows, k)), dev(act_rows(rng, rows, k))
    w = dev((1.0 + 0.1 * rng.standard_normal(k)).astype(np.float32))
    for al in (False, True):
        nat = (lambda y: cu.q8_K_unpitch(y)) if al else (lambda y: y)
        assert torch.equal(nat(cu.prep_q8_K(x, aligned=al)), cu.quantize_q8_K(x))
        for off in (False, True):
            assert torch.equal(nat(cu.prep_q8_K(x, 1, w, 1e-5, off, aligned=al)), cu.quantize_q8_K(cu.rmsnorm(x, w, 1e-5, off)))
        assert torch.equal(nat(cu.prep_q8_K(x, 2, u, aligned=al)), cu.quantize_q8_K(cu.silu_mul(x, u)))
        assert torch.equal(nat(cu.prep_q8_K(x, 2, u, pdl=True, aligned=al)), cu.quantize_q8_K(cu.silu_mul(x, u)))


@pytest.mark.parametrize("M,Ns,K", [(1, (64,), 256), (1, (4096, 1024, 1024), 4096), (1, (4096,), 11008), (2, (50, 33), 11008), (3, (7,), 1280),
                                    (1, (11008, 11008), 4096), (1, (1,), 8192), (4, (3, 1, 2), 4608),
                                    (5, (130, 2, 61), 4352), (8, (33, 40), 1024), (8, (256,), 4096)])
def test_gemv_q4k_segments(cu, oracle, M, Ns, K):
    rng = np.random.default_rng(M * 7 + sum(Ns) + K)
    Wn = [rand_q4k(rng, N, K) for N in Ns]
    bn = [(rng.standard_normal(N) * 0.02).astype(np.float32) if i != 1 else None for i, N in enumerate(Ns)]
    xn = act_rows(rng, M, K)
    Ws, bs, x = [dev(W) for W in Wn], [dev(b) if b is not None else None for b in bn], dev(xn)
    want = [cu.linear(W, x, b, wtype=cu.TYPE_Q4_K).cpu().numpy() for W, b in zip(Ws, bs)]
    if sum(Ns) * K <= 1 << 24:  # and against the oracle itself where it finishes in seconds
        for W, b, w in zip(Wn, bn, want):
            kq_check(w, oracle.linear_kq(W, xn, b if b is not None else np.zeros(W.shape[0], np.float32), cu.TYPE_Q4_K))
    for al in (False, True):
        xq = cu.prep_q8_K(x, aligned=al)
        for pdl, no_areg in ((False, False), (True, False), (True, True)):  # one row, K <= 4096, 304-byte blocks: activations held in registers
            got = cu.gemv_q4k(Ws, xq, K, bs, pdl=pdl, aligned=al, no_areg=no_areg)
            for g, w in zip(got, want):
                kq_check(g.cpu().numpy(), w)
        if len(Ns) == 1:
            res = rng.standard_normal((M, Ns[0])).astype(np.float32)
            got = cu.gemv_q4k(Ws, xq, K, bs, residual=dev(res), pdl=True, aligned=al)[0].cpu().numpy()
            assert np.abs(got - (res + want[0])).max() <= 2e-5 * np.abs(want[0]).max() + 1e-6


def test_q4k_decode_chain_with_pdl_matches_op_by_op(cu):
    """One MLP half of the step as the plan launches it (every kernel a programmatic dependent of the one before) against the separate ops."""
    rng = np.random.default_rng(11)
    H, F, M = 1024, 2816, 2
    gate, up, down = dev(rand_q4k(rng, F, H)), dev(rand_q4k(rng, F, H)), dev(rand_q4k(rng, H, F))
    w = dev((1.0 + 0.1 * rng.standard_normal(H)).astype(np.float32))
    x = dev((rng.standard_normal((M, H)) * 0.05).astype(np.float32))
    t = cu.rmsnorm(x, w, 1e-6)
    g = cu.silu_mul(cu.linear(gate, t, wtype=cu.TYPE_Q4_K), cu.linear(up, t, wtype=cu.TYPE_Q4_K))
    want = cu.add(x, cu.linear(down, g, wtype=cu.TYPE_Q4_K)).cpu().numpy()
    first = None
    for _ in range(3):
        xq = cu.prep_q8_K(x, 1, w, 1e-6, pdl=True, aligned=True)
        gg, uu = cu.gemv_q4k([gate, up], xq, H, pdl=True, aligned=True)
        gq = cu.prep_q8_K(gg, 2, uu, pdl=True, aligned=True)
        got = cu.gemv_q4k([down], gq, F, residual=x, pdl=True, aligned=True)[0].cpu().numpy()
        # the second Linear requantises the first one's output: a rounding flip there moves a value by one Q8_K step
        assert np.abs(got - want).max() <= 2e-3 * np.abs(want).max()
        assert first is None or np.array_equal(got, first)  # and the chain itself is deterministic
        first = got


def test_linear_q4_K_rejects_k_not_multiple_of_256(cu):
    import torch
    with pytest.raises(cu.MllmcuError):
        cu.linear(torch.zeros((4, 144), dtype=torch.uint8, device="cuda"), torch.zeros((1, 160), device="cuda"), wtype=cu.TYPE_Q4_K)


# ------------------------------------------------------------------ a4: Q4_0_4_4
def test_q4_0_4x4_deinterleave_golden_and_linear(cu, oracle, g2):
    rows = cu.q4_0_4x4_to_q4_0(dev(g2["q44_packed"]), 8, 128).cpu().numpy()
    assert np.array_equal(rows, g2["q44_rows"])
    y = cu.linear(dev(rows), dev(g2["q44_x"][None, :])).cpu().numpy()[0]
    assert np.abs(y - g2["q44_y"]).max() <= 2e-5 * np.abs(g2["q44_y"]).max()  # vs the reference's gemv_q4_0_4x4_q8_0


def test_q4_0_4x4_deinterleave_large(cu, oracle):
    rng = np.random.default_rng(11)
    N, K = 256, 4096
    rows = rng.integers(0, 256, (N, K // 32 * 18), dtype=np.uint8)
    packed = oracle.only.q4_0_4x4_pack(rows, K)
    assert np.array_equal(cu.q4_0_4x4_to_q4_0(dev(packed), N, K).cpu().numpy(), rows)


# ------------------------------------------------------------------ a7: eager softmax / matmul
def test_softmax_golden_bit_exact(cu, g2):
    assert np.array_equal(bits(cu.softmax(dev(g2["sm_x"])).cpu().numpy()), bits(g2["sm_y"]))
    got = cu.softmax(dev(g2["sm_x"][:, :, :8].copy()), causal=True, old_dim=3).cpu().numpy()
    assert np.array_equal(bits(got), bits(g2["sm_y_causal"]))


@pytest.mark.parametrize("outer,S,n,causal,old", [(32, 1, 641, False, 0), (32, 1, 640, False, 0), (8, 16, 16, True, 0), (4, 5, 37, True, 32),
                                                  (16, 577, 577, False, 0), (16, 576, 576, False, 0)])
def test_softmax_vs_oracle(cu, oracle, outer, S, n, causal, old):
    rng = np.random.default_rng(outer + S + n)
    x = (rng.standard_normal((outer, S, n)) * 4).astype(np.float32)
    got = cu.softmax(dev(x), causal, old).cpu().numpy()
    want = oracle.softmax(x, causal, old)
    # rows whose visible length is a multiple of 8 run entirely through the polynomial exp in the reference's lane order: bit-identical.
    # Other rows have <= 7 tail elements through libm expf (0.502 ulp) vs a correctly rounded exp here: <= 1 ulp on those, which
    # can move the row sum by an ulp.
    vis = np.full((S,), n) if not (causal and S > 1) else np.minimum(np.arange(S) + 1 + old, n)
    exact = vis % 8 == 0
    assert np.array_equal(bits(got[:, exact]), bits(want[:, exact]))
    ulp = np.abs(bits(got).astype(np.int64) - bits(want).astype(np.int64))
    assert ulp.max() <= 4


def test_eager_attention_equals_flash(cu, oracle):
    """q/sqrt(D) -> QK^T -> causal softmax -> PV (the eager spelling, modeling_transformer.hpp:187-197) against the flash kernel"""
    import torch
    rng = np.random.default_rng(21)
    Sq, Sk, H, D = 8, 24, 4, 64
    q = rng.standard_normal((H, Sq, D)).astype(np.float32)
    k = rng.standard_normal((H, Sk, D)).astype(np.float16).astype(np.float32)
    v = rng.standard_normal((H, Sk, D)).astype(np.float16).astype(np.float32)
    qs = cu.div_scalar(dev(q), float(np.sqrt(np.float32(D))))
    assert np.array_equal(bits(qs.cpu().numpy()), bits(q / np.sqrt(np.float32(D))))
    sc = cu.sgemm(qs, dev(k))
    p = cu.softmax(sc, causal=True, old_dim=Sk - Sq)
    o = cu.sgemm(p, dev(v), b_is_kn=True).cpu().numpy()
    qf = dev(np.ascontiguousarray(q.transpose(1, 0, 2))[None])
    kf = dev(np.ascontiguousarray(k.transpose(1, 0, 2))[None]).half()
    vf = dev(np.ascontiguousarray(v.transpose(1, 0, 2))[None]).half()
    of = cu.flash_attn_f16kv(qf, kf, vf, causal=True).cpu().numpy()[0].transpose(1, 0, 2)
    assert np.abs(o - of).max() <= 1e-5


def test_sgemm_vs_numpy(cu):
    rng = np.random.default_rng(2)
    A = rng.standard_normal((3, 70, 130)).astype(np.float32)
    B = rng.standard_normal((3, 45, 130)).astype(np.float32)
    bias = rng.standard_normal(45).astype(np.float32)
    got = cu.sgemm(dev(A), dev(B), dev(bias)).cpu().numpy()
    want = np.einsum("bmk,bnk->bmn", A.astype(np.float64), B.astype(np.float64)) + bias
    assert np.abs(got - want).max() <= 1e-5 * np.abs(want).max()


# ------------------------------------------------------------------ a15: GELU tables
@pytest.mark.parametrize("kind,tn,yn", [(0, "gelu_table", "gelu_y"), (1, "gelu_quick_table", "gelu_quick_y")])
def test_gelu_table_bit_exact(cu, oracle, g2, kind, tn, yn):
    t = dev(g2[tn].view(np.int16))
    assert np.array_equal(bits(cu.gelu_table(dev(g2["gelu_x"]), t).cpu().numpy()), bits(g2[yn]))
    rng = np.random.default_rng(kind)
    x = (rng.standard_normal(577 * 4096 // 8 + 3) * 3).astype(np.float32)
    assert np.array_equal(bits(cu.gelu_table(dev(x), t).cpu().numpy()), bits(oracle.only.gelu(g2[tn], x)))


# ------------------------------------------------------------------ a10 / a11: M-RoPE, vision RoPE
def test_mrope_bit_exact(cu, oracle):
    rng = np.random.default_rng(4)
    S, H, D = 37, 12, 128
    pos = np.stack([np.arange(S) // 3, rng.integers(0, 40, S), rng.integers(0, 40, S)]).astype(np.float32)
    s, c = oracle.only.mrope_tables(pos, D, 1e6, [16, 24, 24, 16, 24, 24])
    x = rng.standard_normal((S, H, D)).astype(np.float32)
    want = oracle.only.rope_apply(4, x, 0, s, c)
    got = cu.rope_mrope(dev(x[None]), dev(s), dev(c)).cpu().numpy()[0]
    assert np.array_equal(bits(got), bits(want))
    want16 = oracle.only.rope_apply(4, x, 0, s, c, out_fp16=True)
    got16 = cu.rope_mrope(dev(x[None]), dev(s), dev(c), out_f16=True).cpu().numpy()[0].view(np.uint16)
    assert np.array_equal(got16, want16)


def test_vision_rope(cu, oracle):
    rng = np.random.default_rng(5)
    S, H, D = 196, 16, 80
    x = rng.standard_normal((S, H, D)).astype(np.float32)
    ang = rng.uniform(0, 74, (S, D // 2)).astype(np.float32)
    want = oracle.only.rope_vision(x, ang)
    # host-built sin/cos with libm (std::sin / std::cos, what the backend op uploads and the reference evaluates inline): the result is
    # bit-identical to the restatement, which is pinned to the compiled CPUVisionRoPEFunc (tests/test_oracle_refops.py)
    import ctypes
    libm = ctypes.CDLL("libm.so.6")
    libm.sinf.restype = libm.cosf.restype = ctypes.c_float
    libm.sinf.argtypes = libm.cosf.argtypes = [ctypes.c_float]
    sn = np.array([libm.sinf(float(a)) for a in ang.ravel()], dtype=np.float32).reshape(ang.shape)
    cs = np.array([libm.cosf(float(a)) for a in ang.ravel()], dtype=np.float32).reshape(ang.shape)
    got = cu.rope_vision(dev(x[None]), sin_t=dev(sn), cos_t=dev(cs)).cpu().numpy()[0]
    assert np.array_equal(bits(got), bits(want))
    # device sin/cos: <= 2 ulp on sin/cos -> 1e-6 absolute on O(1) data
    got = cu.rope_vision(dev(x[None]), angles=dev(ang)).cpu().numpy()[0]
    assert np.abs(got - want).max() <= 2e-6 * np.abs(x).max()


# ------------------------------------------------------------------ a17: patch embed
def test_conv2d_golden(cu, g2):
    y = cu.conv2d_valid(dev(g2["conv_x"]), dev(g2["conv_w"]), dev(g2["conv_b"]), (14, 14)).cpu().numpy()
    assert np.abs(y - g2["conv_y"]).max() <= 1e-5 * np.abs(g2["conv_y"]).max()
    y = cu.conv2d_valid(dev(g2["conv_x"]), dev(g2["conv_w"]), None, (7, 7)).cpu().numpy()
    assert np.abs(y - g2["conv_y_s7"]).max() <= 1e-5 * np.abs(g2["conv_y_s7"]).max()


def test_clip_patch_embed_shape(cu, oracle):
    """CLIP ViT-L/14-336: 3x336x336, k = 14 -> 576 patches x 1024, channels-last output = the [patches, hidden] the tower uses"""
    rng = np.random.default_rng(6)
    x = rng.standard_normal((1, 3, 336, 336)).astype(np.float32)
    w = (rng.standard_normal((1024, 3, 14, 14)) * 0.02).astype(np.float32)
    want = oracle.conv2d_valid(x, w, None, (14, 14))
    got = cu.conv2d_valid(dev(x), dev(w), None, (14, 14)).cpu().numpy()
    assert np.abs(got - want).max() <= 2e-5 * np.abs(want).max()
    got_cl = cu.conv2d_valid(dev(x), dev(w), None, (14, 14), channels_last=True).cpu().numpy()
    assert np.array_equal(got_cl[0], got[0].reshape(1024, 576).T)


def test_conv3d_qwen2vl_patch_embed(cu, oracle):
    rng = np.random.default_rng(7)
    x = rng.standard_normal((1, 3, 2, 56, 70)).astype(np.float32)
    w = (rng.standard_normal((96, 3, 2, 14, 14)) * 0.02).astype(np.float32)
    b = rng.standard_normal(96).astype(np.float32)
    for cl in (False, True):
        want = oracle.only.conv3d_valid(x, w, b, (2, 14, 14), channels_last=cl)
        got = cu.conv3d_valid(dev(x), dev(w), dev(b), (2, 14, 14), channels_last=cl).cpu().numpy()
        assert got.shape == want.shape and np.abs(got - want).max() <= 2e-5 * np.abs(want).max()


def test_conv3d_one_patch_per_batch_entry(cu, oracle):
    """How modeling_qwen2_vl.hpp calls it: the processor hands over [patches, 3, 2, 14, 14] and the kernel covers a whole entry -- one output
    position per entry (sgemm.cu runs it as one plain product over the entries)."""
    rng = np.random.default_rng(8)
    x = rng.standard_normal((150, 3, 2, 14, 14)).astype(np.float32)
    w = (rng.standard_normal((96, 3, 2, 14, 14)) * 0.02).astype(np.float32)
    b = rng.standard_normal(96).astype(np.float32)
    for cl in (False, True):
        want = oracle.only.conv3d_valid(x, w, b, (2, 14, 14), channels_last=cl)
        got = cu.conv3d_valid(dev(x), dev(w), dev(b), (2, 14, 14), channels_last=cl).cpu().numpy()
        assert got.shape == want.shape and np.abs(got - want).max() <= 2e-5 * np.abs(want).max()


# ------------------------------------------------------------------ ViT attention (fp32 K/V, non-causal)
# default workspace: K / V split once into fp16 hi / lo planes, streamed with cp.async; small_ws: every CTA splits the tiles it reads.
# Same split values, same MMA order: the two agree bit for bit, and both hold the 2e-5 bar against float64.
@pytest.mark.parametrize("S,H,D", [(577, 16, 64), (100, 16, 80), (64, 4, 128), (1111, 3, 80), (130, 2, 128)])
def test_flash_attn_f32kv_noncausal(cu, S, H, D):
    import torch
    rng = np.random.default_rng(S + D)
    q, k, v = (rng.standard_normal((1, S, H, D)).astype(np.float32) for _ in range(3))
    got_t = cu.flash_attn_f32kv(dev(q), dev(k), dev(v), causal=False)
    assert torch.equal(got_t, cu.flash_attn_f32kv(dev(q), dev(k), dev(v), causal=False, small_ws=True))
    got = got_t.cpu().numpy()[0]
    s = np.einsum("qhd,khd->hqk", q[0].astype(np.float64), k[0].astype(np.float64)) / np.sqrt(D)
    p = np.exp(s - s.max(-1, keepdims=True))
    p /= p.sum(-1, keepdims=True)
    want = np.einsum("hqk,khd->qhd", p, v[0].astype(np.float64))
    assert np.abs(got - want).max() <= 2e-5


def test_flash_attn_f32kv_batched_and_causal_planes(cu):
    """two batch entries with a strided (non-dense) K/V batch stride, causal, GQA: the planes are dense whatever the source stride"""
    import torch
    rng = np.random.default_rng(5)
    B, S, Hq, Hkv, D = 2, 200, 4, 2, 64
    q = dev(rng.standard_normal((B, S, Hq, D)).astype(np.float32))
    kbig, vbig = (dev(rng.standard_normal((B, S + 56, Hkv, D)).astype(np.float32)) for _ in range(2))
    k, v = kbig[:, :S], vbig[:, :S]
    a = cu.flash_attn_f32kv(q, k, v, causal=True)
    b = cu.flash_attn_f32kv(q, k, v, causal=True, small_ws=True)
    c = cu.flash_attn_f32kv(q, k.contiguous(), v.contiguous(), causal=True)
    assert torch.equal(a, b) and torch.equal(a, c)


def test_flash_attn_f32kv_matches_reference_variant(cu, ref_lib):
    if ref_lib is None:
        pytest.skip("oracle/_ref not built here")
    import ctypes as C
    rng = np.random.default_rng(9)
    Sq = Sk = 64
    H, D = 4, 64
    q, k, v = (rng.standard_normal((Sq, H, D)).astype(np.float32) for _ in range(3))
    o = np.empty_like(q)
    p = lambda a: a.ctypes.data_as(C.c_void_p)
    ref_lib.lib.ref_flash_attn_f32kv(p(q), p(k), p(v), p(o), Sq, Sk, H, H, D, 0, 1)
    got = cu.flash_attn_f32kv(dev(q[None]), dev(k[None]), dev(v[None]), causal=False).cpu().numpy()[0]
    assert np.abs(got - o).max() <= 1e-5


# ------------------------------------------------------------------ a18: strided copy
def test_copy_strided_transpose_and_cat(cu):
    import torch
    rng = np.random.default_rng(8)
    B, S, H, D = 2, 9, 4, 16
    x = rng.standard_normal((B, S, H, D)).astype(np.float32)
    # BSHD -> BHSD physical transpose (CPUtransposeFunction HEAD<->SEQUENCE)
    dst = torch.empty((B, H, S, D), device="cuda")
    cu.copy_strided(dev(x), dst, (B, H, S, D), (S * H * D, D, H * D, 1), (H * S * D, S * D, D, 1))
    assert np.array_equal(dst.cpu().numpy(), x.transpose(0, 2, 1, 3))
    # cat along SEQUENCE: second tensor lands behind the first
    y = rng.standard_normal((B, 5, H, D)).astype(np.float32)
    out = torch.zeros((B, S + 5, H, D), device="cuda")
    cu.copy_strided(dev(x), out, (B, S, H, D), (S * H * D, H * D, D, 1), ((S + 5) * H * D, H * D, D, 1))
    cu.copy_strided(dev(y), out[:, S:], (B, 5, H, D), (5 * H * D, H * D, D, 1), ((S + 5) * H * D, H * D, D, 1))
    assert np.array_equal(out.cpu().numpy(), np.concatenate([x, y], axis=1))
    # fp16 KV head replication (n_rep = 2)
    kv = torch.from_numpy(rng.standard_normal((1, 6, 2, 8)).astype(np.float16)).cuda()
    rep = torch.empty((1, 6, 4, 8), dtype=torch.float16, device="cuda")
    for r in range(2):
        cu.copy_strided(kv, rep[:, :, r:], (1, 6, 2, 8), (96, 16, 8, 1), (192, 32, 16, 1))
    assert np.array_equal(rep.cpu().numpy(), np.repeat(kv.cpu().numpy(), 2, axis=2))
