"""GPU parity tests: every C-ABI kernel (include/mllmcu.h, called through ctypes) against
the CPU oracle on the same seeded inputs.  Bit-exact for bytes / integers; stated
tolerances for fp32."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def bits(a):
    return np.ascontiguousarray(a, dtype=np.float32).view(np.uint32)


def dev(a):
    import torch
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


def act_rows(rng, rows, k):
    x = (rng.standard_normal((rows, k)) * rng.uniform(0.01, 8, (rows, 1))).astype(np.float32)
    if rows > 2 and k >= 64:
        x[1, :32] = 0.0
        x[2, :32] = np.arange(32, dtype=np.float32) * 0.5
        x[2, 31] = 63.5  # ties-to-even
    return x


def test_q8_0_blocks_bit_exact_golden(cu, golden):
    y = cu.quantize_q8_0(dev(golden["q8_x"])).cpu().numpy()
    assert np.array_equal(y, golden["q8_blocks"])


@pytest.mark.parametrize("rows,k", [(1, 32), (1, 4096), (5, 11008), (64, 2048), (3, 5632)])
def test_q8_0_blocks_bit_exact(cu, oracle, rows, k):
    rng = np.random.default_rng(rows * 131 + k)
    x = act_rows(rng, rows, k)
    assert np.array_equal(cu.quantize_q8_0(dev(x)).cpu().numpy(), oracle.quantize_q8_0(x))


def test_q8_0_rejects_ragged_k(cu):
    import torch
    x = torch.zeros((1, 48), device="cuda")
    with pytest.raises(cu.MllmcuError):
        cu.quantize_q8_0(x)


@pytest.mark.parametrize("M,N,K", [(1, 64, 256), (3, 33, 4096), (16, 40, 512)])
def test_integer_block_dots_bit_exact(cu, oracle, M, N, K):
    rng = np.random.default_rng(M + N + K)
    W = oracle.quantize_q4_0((rng.standard_normal((N, K)) * 0.02).astype(np.float32))
    x = act_rows(rng, M, K)
    xq = oracle.quantize_q8_0(x)
    got = cu.block_isums_q4_0_q8_0(dev(W), dev(xq), K).cpu().numpy()
    for n in range(0, N, max(1, N // 5)):
        for m in range(M):
            assert np.array_equal(got[n, m], oracle.only.block_isums_q4_0_q8_0(W[n], xq[m], K))


def lin_check(y, want):
    # fp32 accumulation order differs from the CPU's 8 AVX lanes: max-abs <= 1e-3 * max|y|
    # (SURVEY 8c bar) -- in practice ~1e-6 relative -- and cosine >= 0.99999.
    y = y.astype(np.float64)
    want = want.astype(np.float64)
    assert np.abs(y - want).max() <= 2e-5 * max(np.abs(want).max(), 1e-6)
    cos = (y * want).sum() / (np.linalg.norm(y) * np.linalg.norm(want) + 1e-30)
    assert cos >= 0.99999


@pytest.mark.parametrize("M,N,K", [(1, 128, 256), (1, 4096, 4096), (2, 300, 2048), (3, 96, 5632), (8, 64, 11008),
                                    (9, 130, 512), (64, 256, 1024), (130, 70, 2048)])
@pytest.mark.parametrize("bias", [False, True])
def test_linear_q4_0(cu, oracle, M, N, K, bias):
    rng = np.random.default_rng(M * 7 + N * 3 + K)
    W = oracle.quantize_q4_0((rng.standard_normal((N, K)) * 0.02).astype(np.float32))
    x = act_rows(rng, M, K)
    b = (rng.standard_normal(N) * 0.02).astype(np.float32) if bias else None
    y = cu.linear(dev(W), dev(x), dev(b) if bias else None).cpu().numpy()
    lin_check(y, oracle.linear(W, x, b))


@pytest.mark.parametrize("M,N,K", [(9, 128, 64), (128, 128, 256), (130, 256, 512), (300, 384, 2048), (512, 128, 11008), (33, 512, 4096)])
@pytest.mark.parametrize("bias", [False, True])
def test_linear_q4_0_tensor_core_path(cu, oracle, M, N, K, bias):
    """Prefill shapes that take the tcgen05 kernel (M > 8, N % 128 == 0, K % 64 == 0): per-block integer sums come out of
    the tensor core exactly (fp16 integers, fp32 accumulate), so the result obeys the same bar as the CUDA-core path."""
    rng = np.random.default_rng(M + N + K)
    W = oracle.quantize_q4_0((rng.standard_normal((N, K)) * 0.02).astype(np.float32))
    x = act_rows(rng, M, K)
    b = (rng.standard_normal(N) * 0.02).astype(np.float32) if bias else None
    y = cu.linear(dev(W), dev(x), dev(b) if bias else None).cpu().numpy()
    lin_check(y, oracle.linear(W, x, b))


def test_linear_tensor_core_block_sums_exact(cu, oracle):
    """With d_w = d_x = 1 the output IS the sum of integer block sums: must equal the integer arithmetic exactly."""
    rng = np.random.default_rng(5)
    M, N, K = 256, 256, 1024
    blocks = rng.integers(0, 256, (N, K // 32, 18), dtype=np.uint8)
    one = np.float16(1.0).view(np.uint16)
    blocks[:, :, 0] = one & 0xFF
    blocks[:, :, 1] = one >> 8
    W = blocks.reshape(N, -1)
    q = rng.integers(-127, 128, (M, K)).astype(np.float32)
    q[:, ::32] = 127.0  # amax = 127 in every block -> d_x = 1, q8 = the integers themselves
    y = cu.linear(dev(W), dev(q)).cpu().numpy()
    qs = blocks[:, :, 2:]
    w = np.concatenate([(qs & 0x0F).astype(np.int64) - 8, (qs >> 4).astype(np.int64) - 8], axis=2).reshape(N, K)
    want = q.astype(np.int64) @ w.T
    assert np.array_equal(y.astype(np.int64), want) and np.all(y == np.round(y))


def _with_env(name, val, fn):
    import os
    old = os.environ.get(name)
    os.environ[name] = val
    try:
        return fn()
    finally:
        if old is None:
            del os.environ[name]
        else:
            os.environ[name] = old


@pytest.mark.parametrize("M,N,K", [(128, 128, 256), (300, 384, 2048), (512, 256, 11008)])
def test_linear_tensor_core_both_kernels(cu, oracle, M, N, K):
    """The per-block integer-sum kernel (default) and the scaled-operand kernel (MLLMCU_TC_EXACT=0: scales folded into exactly
    split fp16 operands, activations in tensor memory, chained fp32 accumulation) against the oracle and against each other."""
    rng = np.random.default_rng(M + N + K + 1)
    W = oracle.quantize_q4_0((rng.standard_normal((N, K)) * 0.02).astype(np.float32))
    x = act_rows(rng, M, K)
    x[0] *= 500.0   # a row of large activations (|a| ~ 4000) and a tiny one: the hi/lo split must hold over the range
    x[3] *= 1e-4
    b = (rng.standard_normal(N) * 0.02).astype(np.float32)
    want = oracle.linear(W, x, b)
    ys = _with_env("MLLMCU_TC_EXACT", "0", lambda: cu.linear(dev(W), dev(x), dev(b)).cpu().numpy())
    ye = _with_env("MLLMCU_TC_EXACT", "1", lambda: cu.linear(dev(W), dev(x), dev(b)).cpu().numpy())
    for m in (0, 3, 7, M - 1):  # per row: each row has its own scale
        lin_check(ys[m], want[m])
        lin_check(ye[m], want[m])
    lin_check(ys, want)
    lin_check(ye, want)
    assert not np.array_equal(ys, ye) or M == 0  # two different kernels really ran


def test_linear_tensor_core_exact_kernel_block_sums(cu):
    rng = np.random.default_rng(6)
    M, N, K = 128, 128, 512
    blocks = rng.integers(0, 256, (N, K // 32, 18), dtype=np.uint8)
    one = np.float16(1.0).view(np.uint16)
    blocks[:, :, 0] = one & 0xFF
    blocks[:, :, 1] = one >> 8
    q = rng.integers(-127, 128, (M, K)).astype(np.float32)
    q[:, ::32] = 127.0
    y = _with_env("MLLMCU_TC_EXACT", "1", lambda: cu.linear(dev(blocks.reshape(N, -1)), dev(q)).cpu().numpy())
    qs = blocks[:, :, 2:]
    w = np.concatenate([(qs & 0x0F).astype(np.int64) - 8, (qs >> 4).astype(np.int64) - 8], axis=2).reshape(N, K)
    assert np.array_equal(y.astype(np.int64), q.astype(np.int64) @ w.T)


def test_linear_tensor_core_fp16_out_strided(cu, oracle):
    import torch
    rng = np.random.default_rng(8)
    M, N, K, ld = 40, 128, 512, 160
    W = oracle.quantize_q4_0((rng.standard_normal((N, K)) * 0.05).astype(np.float32))
    x = act_rows(rng, M, K)
    out = torch.zeros((M + 2, ld), dtype=torch.float16, device="cuda")
    cu.linear(dev(W), dev(x), out_f16=True, out=out[1:], ldy=ld)
    got = out.cpu().numpy()
    want = oracle.linear(W, x, out_fp16=True).view(np.float16).astype(np.float32)
    assert np.all(got[0] == 0) and np.all(got[M + 1] == 0) and np.all(got[:, N:] == 0)
    assert np.abs(got[1:M + 1, :N].astype(np.float32) - want).max() <= 1e-3 * np.abs(want).max()


def test_linear_q4_0_golden(cu, golden):
    y = cu.linear(dev(golden["q4_blocks"]), dev(golden["q8_x"]), dev(golden["lin_bias"])).cpu().numpy()
    lin_check(y, golden["lin_y_bias"])


@pytest.mark.parametrize("M", [1, 20])
def test_linear_q8_0_weights(cu, oracle, M):
    from oracle import pyoracle as po
    rng = np.random.default_rng(M)
    N, K = 72, 1024
    W = oracle.quantize_q8_0((rng.standard_normal((N, K)) * 0.02).astype(np.float32))
    x = act_rows(rng, M, K)
    y = cu.linear(dev(W), dev(x), wtype=cu.TYPE_Q8_0).cpu().numpy()
    lin_check(y, oracle.linear(W, x, wtype=po.Q8_0))


def test_linear_fp16_into_strided_slice(cu, oracle):
    """v_proj writes fp16 straight into a KV-cache slice (row stride = cache row)."""
    import torch
    rng = np.random.default_rng(3)
    M, N, K, ld = 5, 64, 512, 64
    W = oracle.quantize_q4_0((rng.standard_normal((N, K)) * 0.05).astype(np.float32))
    x = act_rows(rng, M, K)
    cache = torch.zeros((16, ld), dtype=torch.float16, device="cuda")
    cu.linear(dev(W), dev(x), out_f16=True, out=cache[4:], ldy=ld)
    got = cache.cpu().numpy()
    want = oracle.linear(W, x, out_fp16=True).view(np.float16)
    assert np.all(got[:4] == 0) and np.all(got[9:] == 0)
    # fp16 of values that agree to ~1e-6 relative: at most 1 fp16 ulp apart
    assert np.abs(got[4:9].astype(np.float32) - want.astype(np.float32)).max() <= 1e-3 * np.abs(want.astype(np.float32)).max()


def test_linear_linearity_full_size(cu, oracle):
    """Full LLaMA-2-7B gate_proj shape (11008 x 4096): y(x) for a one-hot-block x equals the
    dequantised weight column sums -- a size-independent property checked without the O(NK) oracle."""
    import torch
    rng = np.random.default_rng(11)
    N, K = 11008, 4096
    Wt = torch.randint(0, 256, (N, K // 32 * 18), dtype=torch.uint8, device="cuda")
    # sane fp16 scales: overwrite d bytes of every block with fp16(0.01)
    Wv = Wt.view(N, K // 32, 18)
    d16 = np.float16(0.01).view(np.uint16)
    Wv[:, :, 0] = int(d16 & 0xFF)
    Wv[:, :, 1] = int(d16 >> 8)
    x = np.zeros((1, K), dtype=np.float32)
    x[0, 64:96] = 127.0  # block 2: q8 = 127 each, d_x = 1.0 exactly
    y = cu.linear(Wt, dev(x)).cpu().numpy()[0]
    blk = Wv[:, 2, 2:].cpu().numpy()
    isum = ((blk & 0x0F).astype(np.int64) - 8).sum(1) + ((blk >> 4).astype(np.int64) - 8).sum(1)
    want = (np.float32(np.float16(0.01)) * np.float32(1.0)) * (isum * 127).astype(np.float32)
    assert np.array_equal(bits(y), bits(want.astype(np.float32)))


@pytest.mark.parametrize("rows,dim", [(1, 4096), (7, 2048), (512, 4096), (3, 100)])
def test_rmsnorm_bit_exact(cu, oracle, rows, dim):
    rng = np.random.default_rng(rows + dim)
    x = (rng.standard_normal((rows, dim)) * 3).astype(np.float32)
    w = (1 + 0.02 * rng.standard_normal(dim)).astype(np.float32)
    y = cu.rmsnorm(dev(x), dev(w), 1e-6).cpu().numpy()
    want = oracle.only.rmsnorm(x, w, 1e-6)
    # double-precision sum of squares makes the float mean order-independent except on
    # exact rounding ties: allow <= 1 ulp on a vanishing fraction, typically bit-exact
    diff = np.abs(bits(y).astype(np.int64) - bits(want).astype(np.int64))
    assert diff.max() <= 1
    assert (diff > 0).mean() < 1e-3


def test_layernorm(cu, oracle):
    rng = np.random.default_rng(2)
    x = rng.standard_normal((9, 1024)).astype(np.float32)
    w = rng.standard_normal(1024).astype(np.float32)
    b = rng.standard_normal(1024).astype(np.float32)
    y = cu.layernorm(dev(x), dev(w), dev(b), 1e-6).cpu().numpy()
    assert np.abs(y - oracle.only.layernorm(x, w, b, 1e-6)).max() <= 2e-5


@pytest.mark.parametrize("rtype", [2, 4])
@pytest.mark.parametrize("f16", [False, True])
def test_rope_bit_exact(cu, oracle, rtype, f16):
    rng = np.random.default_rng(rtype)
    B, S, H, D, pos0 = 2, 5, 4, 128, 9
    s, c = oracle.only.rope_tables(rtype, 64, D, 10000.0)
    x = rng.standard_normal((B, S, H, D)).astype(np.float32)
    y = cu.rope(rtype, dev(x), pos0, dev(s), dev(c), out_f16=f16).cpu().numpy()
    for b in range(B):
        want = oracle.only.rope_apply(rtype, x[b], pos0, s, c, out_fp16=f16)
        if f16:
            assert np.array_equal(y[b].view(np.uint16), want)
        else:
            assert np.array_equal(bits(y[b]), bits(want))


@pytest.mark.parametrize("f16", [False, True])
def test_rope_llama_prefill_shape_bit_exact(cu, oracle, f16):
    """A prefill-sized LLaMA-type rotation takes the 16-byte kernel (two pairs per thread): same bits as the oracle."""
    rng = np.random.default_rng(12)
    B, S, H, D, pos0 = 1, 40, 8, 128, 3
    s, c = oracle.only.rope_tables(2, 64, D, 10000.0)
    x = rng.standard_normal((B, S, H, D)).astype(np.float32)
    y = cu.rope(2, dev(x), pos0, dev(s), dev(c), out_f16=f16).cpu().numpy()
    want = oracle.only.rope_apply(2, x[0], pos0, s, c, out_fp16=f16)
    if f16:
        assert np.array_equal(y[0].view(np.uint16), want)
    else:
        assert np.array_equal(bits(y[0]), bits(want))


def test_rope_partial_and_cache_slice(cu, oracle):
    import torch
    rng = np.random.default_rng(8)
    B, S, H, D, rot, pos0, limit = 1, 3, 2, 64, 32, 4, 16
    s, c = oracle.only.rope_tables(4, 32, rot, 10000.0)
    x = rng.standard_normal((B, S, H, D)).astype(np.float32)
    cache = torch.zeros((B, limit, H, D), dtype=torch.float16, device="cuda")
    cu.rope(4, dev(x), pos0, dev(s), dev(c), rot=rot, out_f16=True, out=cache[:, pos0:], out_bstride=limit * H * D, out_sstride=H * D)
    got = cache.cpu().numpy()
    want = oracle.only.rope_apply(4, x[0], pos0, s, c, rot=rot, out_fp16=True)
    assert np.array_equal(got[0, pos0:pos0 + S].view(np.uint16), want)
    assert np.all(got[0, :pos0] == 0) and np.all(got[0, pos0 + S:] == 0)


def test_silu_bit_exact(cu, oracle, golden):
    y = cu.silu(dev(golden["silu_x"])).cpu().numpy()
    assert np.array_equal(bits(y), bits(golden["silu_y"]))
    rng = np.random.default_rng(4)
    z = (rng.standard_normal(11008 * 3) * 6).astype(np.float32)
    assert np.array_equal(bits(cu.silu(dev(z)).cpu().numpy()), bits(oracle.silu(z)))
    u = rng.standard_normal(z.size).astype(np.float32)
    assert np.array_equal(bits(cu.silu_mul(dev(z), dev(u)).cpu().numpy()), bits(oracle.silu(z) * u))


def test_add_mul_bit_exact(cu):
    rng = np.random.default_rng(5)
    a = rng.standard_normal(4099 * 4).astype(np.float32)
    b = rng.standard_normal(4099 * 4).astype(np.float32)
    assert np.array_equal(bits(cu.add(dev(a), dev(b)).cpu().numpy()), bits(a + b))
    assert np.array_equal(bits(cu.mul(dev(a), dev(b)).cpu().numpy()), bits(a * b))


def test_embedding_q4_0_bit_exact(cu, oracle):
    rng = np.random.default_rng(6)
    V, Hd = 500, 2048
    table = oracle.quantize_q4_0((rng.standard_normal((V, Hd)) * 0.02).astype(np.float32))
    ids = np.array([0, 499, 17, -1, 250], dtype=np.float32)
    got = cu.embedding_q4_0(dev(table), dev(ids), Hd).cpu().numpy()
    assert np.array_equal(bits(got), bits(oracle.only.embedding_q4_0(table, ids, Hd)))


@pytest.mark.parametrize("Sq,Sk,Hq,Hkv,D", [(1, 1, 4, 4, 64), (1, 37, 8, 2, 64), (1, 700, 32, 32, 128), (16, 16, 4, 4, 128),
                                              (4, 12, 4, 2, 64), (7, 19, 4, 4, 64), (33, 33, 2, 1, 128), (128, 128, 8, 8, 64)])
def test_flash_attention(cu, oracle, Sq, Sk, Hq, Hkv, D):
    rng = np.random.default_rng(Sq * 1000 + Sk)
    q = rng.standard_normal((1, Sq, Hq, D)).astype(np.float32)
    k = rng.standard_normal((1, Sk, Hkv, D)).astype(np.float16)
    v = rng.standard_normal((1, Sk, Hkv, D)).astype(np.float16)
    o = cu.flash_attn_f16kv(dev(q), dev(k), dev(v)).cpu().numpy()
    want = oracle.flash_attn_f16kv(q[0], k[0].view(np.uint16), v[0].view(np.uint16))
    # online-softmax tile order differs (per-key vs 4-key tiles): 1e-5 absolute on O(1) values
    assert np.abs(o[0] - want).max() <= 1e-5 * max(1.0, np.abs(want).max())


def test_flash_attention_golden_and_batch(cu, golden):
    import torch
    for tag in ("dec", "pre", "app"):
        q, k, v = golden[f"fa_{tag}_q"], golden[f"fa_{tag}_k"], golden[f"fa_{tag}_v"]
        o = cu.flash_attn_f16kv(dev(q[None]), dev(k.view(np.float16)[None]), dev(v.view(np.float16)[None])).cpu().numpy()
        assert np.abs(o[0] - golden[f"fa_{tag}_o"]).max() <= 1e-5 * max(1.0, np.abs(golden[f"fa_{tag}_o"]).max())
    # batch of 2 reading a cache with a larger batch stride
    q, k, v = golden["fa_dec_q"], golden["fa_dec_k"].view(np.float16), golden["fa_dec_v"].view(np.float16)
    Sk, Hkv, D = k.shape
    limit = 64
    kc = torch.zeros((2, limit, Hkv, D), dtype=torch.float16, device="cuda")
    vc = torch.zeros_like(kc)
    kc[:, :Sk] = dev(k)
    vc[:, :Sk] = dev(v)
    qq = dev(np.stack([q, q]))
    o = cu.flash_attn_f16kv(qq, kc[:, :Sk], vc[:, :Sk], kv_bstride=limit * Hkv * D).cpu().numpy()
    assert np.abs(o[1] - golden["fa_dec_o"]).max() <= 1e-5


def test_argmax_first_max_wins(cu):
    import torch
    x = torch.randn(3, 32000, device="cuda")
    x[0, 777] = 100.0
    x[0, 31999] = 100.0
    x[1, 0] = 50.0
    x[2, 31999] = 60.0
    assert cu.argmax(x).cpu().tolist() == [777, 0, 31999]


# ---------------------------------------------------------------- fused decode Linear (TMA-staged GEMV)
def _q4(oracle, rng, n, k, scale=0.02):
    return oracle.quantize_q4_0((rng.standard_normal((n, k)) * scale).astype(np.float32))


@pytest.mark.parametrize("M,N,K", [(1, 4096, 4096), (1, 64, 256), (2, 1000, 2048), (8, 72, 1024), (3, 4096, 11008), (1, 136, 5504),
                                    (3, 136, 5504), (8, 2048, 5504)])  # K = 5504 (Qwen1.5-1.8B ffn): rows of 3096 B go in pairs at every batch
def test_gemv_fused_plain_bit_equal_to_linear(cu, oracle, M, N, K):
    """Same arithmetic, same order as mllmcu_linear's GEMV: bit-identical; and within tolerance of the oracle."""
    rng = np.random.default_rng(N + K + M)
    W = _q4(oracle, rng, N, K)
    x = act_rows(rng, M, K)
    b = (rng.standard_normal(N) * 0.02).astype(np.float32)
    y = cu.gemv_fused([dev(W)], dev(x), biases=[dev(b)]).cpu().numpy()
    y_lin = cu.linear(dev(W), dev(x), dev(b)).cpu().numpy()
    assert np.array_equal(bits(y), bits(y_lin))
    lin_check(y, oracle.linear(W, x, b))


@pytest.mark.parametrize("M", [1, 4])
def test_gemv_fused_rmsnorm_qkv(cu, oracle, M):
    """RMSNorm -> q/k/v projections in one launch == rmsnorm kernel then three Linears."""
    rng = np.random.default_rng(M)
    K, Nq, Nkv = 2048, 2048, 256
    Ws = [_q4(oracle, rng, n, K) for n in (Nq, Nkv, Nkv)]
    bs = [(rng.standard_normal(n) * 0.02).astype(np.float32) for n in (Nq, Nkv, Nkv)]
    x = (rng.standard_normal((M, K)) * 2).astype(np.float32)
    w = (1 + 0.02 * rng.standard_normal(K)).astype(np.float32)
    y = cu.gemv_fused([dev(v) for v in Ws], dev(x), norm_w=dev(w), eps=1e-5, biases=[dev(v) for v in bs]).cpu().numpy()
    xn = cu.rmsnorm(dev(x), dev(w), 1e-5)
    want = np.concatenate([cu.linear(dev(Wi), xn, dev(bi)).cpu().numpy() for Wi, bi in zip(Ws, bs)], axis=1)
    # the fused prologue sums squares in compensated fp32, the stand-alone kernel in double: same float mean
    assert np.array_equal(bits(y), bits(want))
    xo = oracle.only.rmsnorm(x, w, 1e-5)
    lin_check(y, np.concatenate([oracle.linear(Wi, xo, bi) for Wi, bi in zip(Ws, bs)], axis=1))


def test_gemv_fused_residual_and_silu_mul(cu, oracle):
    rng = np.random.default_rng(12)
    M, K, F = 2, 1024, 2816
    x = (rng.standard_normal((M, K))).astype(np.float32)
    w = (1 + 0.02 * rng.standard_normal(K)).astype(np.float32)
    Wg, Wu, Wd = _q4(oracle, rng, F, K), _q4(oracle, rng, F, K), _q4(oracle, rng, K, F)
    act = cu.gemv_fused([dev(Wg), dev(Wu)], dev(x), norm_w=dev(w), eps=1e-6, epilogue=2)
    xn = cu.rmsnorm(dev(x), dev(w), 1e-6)
    want_act = cu.silu_mul(cu.linear(dev(Wg), xn), cu.linear(dev(Wu), xn))
    assert np.array_equal(bits(act.cpu().numpy()), bits(want_act.cpu().numpy()))
    out = cu.gemv_fused([dev(Wd)], act, epilogue=1, residual=dev(x)).cpu().numpy()
    want = cu.add(cu.linear(dev(Wd), want_act), dev(x)).cpu().numpy()
    assert np.array_equal(bits(out), bits(want))
    xo = oracle.only.rmsnorm(x, w, 1e-6)
    a_o = oracle.silu(oracle.linear(Wg, xo)) * oracle.linear(Wu, xo)
    lin_check(out, oracle.linear(Wd, a_o) + x)


@pytest.mark.parametrize("M", [3, 8])
def test_gemv_fused_residual_batched_odd_row_bytes(cu, oracle, M):
    """Qwen1.5-1.8B down projection at batch 3 / 8 (demo_qwen_batch shapes): K = 5504, residual epilogue."""
    rng = np.random.default_rng(40 + M)
    N, K = 2048, 5504
    W = _q4(oracle, rng, N, K)
    x = act_rows(rng, M, K)
    r = rng.standard_normal((M, N)).astype(np.float32)
    out = cu.gemv_fused([dev(W)], dev(x), epilogue=1, residual=dev(r)).cpu().numpy()
    want = oracle.linear(W, x) + r
    assert np.abs(out - want).max() <= 2e-5 * np.abs(want).max()


def test_gemv_fused_rejects_unsupported(cu, oracle):
    import torch
    rng = np.random.default_rng(1)
    W = dev(_q4(oracle, rng, 32, 96))
    with pytest.raises(cu.MllmcuError):
        cu.gemv_fused([W], torch.zeros((1, 96), device="cuda"))  # K % 128 != 0 -> MLLMCU_ENOSUP


def test_fast_division_is_ieee_exhaustive(cu):
    """Every non-negative fp32 bit pattern (0 .. +inf): the fused kernels' Q8_0 scale arithmetic == IEEE division."""
    import torch
    bad = torch.zeros(1, dtype=torch.int64, device="cuda")
    cu.check(cu.lib().mllmcu_selftest_div(0, 0x7F800000, bad.data_ptr(), cu._stream()))
    torch.cuda.synchronize()
    assert int(bad.item()) == 0


# ---------------------------------------------------------------- argument errors surface as codes, never as silent fallbacks
def test_c_abi_rejects_bad_arguments(cu):
    import torch
    L = cu.lib()
    z = torch.zeros(4096, device="cuda")
    st = cu._stream()
    p = cu._ptr
    # Linear: empty batch, K not a multiple of 32, unknown weight dtype, workspace too small
    assert L.mllmcu_linear(2, p(z), p(z), None, p(z), 0, 64, 256, 256, 64, 0, p(z), z.numel() * 4, st) == -1
    assert L.mllmcu_linear(2, p(z), p(z), None, p(z), 1, 64, 250, 250, 64, 0, p(z), z.numel() * 4, st) == -1
    assert L.mllmcu_linear(3, p(z), p(z), None, p(z), 1, 64, 256, 256, 64, 0, p(z), z.numel() * 4, st) == -2
    assert L.mllmcu_linear(2, p(z), p(z), None, p(z), 1, 64, 256, 256, 64, 0, p(z), 16, st) == -3
    # attention: causal with more queries than keys, heads not a multiple of kv heads, head dim outside the kernels
    assert L.mllmcu_flash_attn_f16kv(p(z), p(z), p(z), p(z), 1, 8, 4, 4, 4, 64, 1, 1024, None, 0, st) == -1
    assert L.mllmcu_flash_attn_f16kv(p(z), p(z), p(z), p(z), 1, 4, 4, 6, 4, 64, 1, 1024, None, 0, st) == -1
    assert L.mllmcu_flash_attn_f16kv(p(z), p(z), p(z), p(z), 1, 4, 4, 4, 4, 48, 1, 1024, None, 0, st) == -2
    # fused decode Linear: batch above 64 (above 8 when N % 16 != 0), K not a multiple of 128
    with pytest.raises(cu.MllmcuError):
        cu.gemv_fused([torch.zeros((64, 144), dtype=torch.uint8, device="cuda")], torch.zeros((65, 256), device="cuda"))
    with pytest.raises(cu.MllmcuError):
        cu.gemv_fused([torch.zeros((72, 144), dtype=torch.uint8, device="cuda")], torch.zeros((9, 256), device="cuda"))  # 9 rows need N % 16 == 0
    torch.cuda.synchronize()  # nothing was launched, nothing is pending


# ---------------------------------------------------------------- full BASELINE sizes: properties instead of the (slow) oracle
@pytest.mark.parametrize("N,K", [(11008, 4096), (4096, 11008), (32000, 4096)])
def test_full_size_decode_and_prefill_linears_agree(cu, N, K):
    """LLaMA-2-7B Linear shapes: the decode kernel (one row, fused TMA GEMV) and the prefill kernel (the same row nine times,
    tcgen05) are independent implementations of the same sum; they must agree row for row, and the nine prefill rows must be
    bit-identical to each other.  Spot rows are also checked against exact integer arithmetic in numpy."""
    import torch
    rng = np.random.default_rng(N + K)
    W = torch.randint(0, 256, (N, K // 32 * 18), dtype=torch.uint8, device="cuda")
    Wv = W.view(N, K // 32, 18)
    Wv[:, :, 1] = 0x1C + torch.randint(0, 4, (N, K // 32), dtype=torch.uint8, device="cuda")  # fp16 scales around 2^-8 .. 2^-5
    x = torch.from_numpy(rng.standard_normal((1, K)).astype(np.float32)).cuda()
    y1 = cu.gemv_fused([W], x).cpu().numpy()[0]
    y9 = cu.linear(W, x.repeat(9, 1).contiguous()).cpu().numpy()
    for m in range(1, 9):
        assert np.array_equal(bits(y9[m]), bits(y9[0]))
    scale = np.abs(y9[0]).max()
    assert np.abs(y1 - y9[0]).max() <= 2e-5 * scale
    # three rows against float64 arithmetic on the dequantised blocks and the Q8_0 activations the kernels use
    xq = cu.quantize_q8_0(x).cpu().numpy().reshape(K // 32, 34)
    dx = xq[:, :2].copy().view(np.float16).astype(np.float64)[:, 0]
    qx = xq[:, 2:].view(np.int8).astype(np.float64)
    Wh = W.cpu().numpy().reshape(N, K // 32, 18)
    for n in (0, N // 2, N - 1):
        dw = Wh[n, :, :2].copy().view(np.float16).astype(np.float64)[:, 0]
        qs = Wh[n, :, 2:]
        qw = np.concatenate([(qs & 0x0F).astype(np.float64) - 8, (qs >> 4).astype(np.float64) - 8], axis=1)
        want = float(np.sum(dw * dx * np.sum(qw * qx, axis=1)))
        assert abs(y1[n] - want) <= 2e-5 * scale and abs(y9[0][n] - want) <= 2e-5 * scale


# ---------------------------------------------------------------- batched decode Linear on the tensor cores (gemv_mma.cu)
@pytest.mark.parametrize("M,Ns,K,epi", [(5, [2048], 2048, 0), (8, [2048, 2048, 2048], 2048, 0), (8, [5504, 5504], 2048, 2), (8, [2048], 5504, 1),
                                        (16, [4096], 4096, 1), (24, [512, 256, 256], 1024, 0), (33, [1024, 1024], 512, 2), (64, [2048], 5504, 1),
                                        (64, [11008, 11008], 4096, 2), (6, [32000], 4096, 0)])
def test_gemv_mma_against_oracle(cu, oracle, M, Ns, K, epi):
    """3..64 activation rows through mllmcu_gemv_fused: RMSNorm prologue (epilogues 0 / 2), bias, residual, silu(gate) * up; every
    output within the Linear tolerance of the oracle (the integer block sums inside are exact: test below)."""
    assert cu.lib().mllmcu_gemv_fused_max_rows() == 64
    rng = np.random.default_rng(M * 7 + K + len(Ns))
    Ws = [_q4(oracle, rng, n, K) for n in Ns]
    x = act_rows(rng, M, K)
    w = (1 + 0.02 * rng.standard_normal(K)).astype(np.float32) if epi != 1 else None
    bs = [(rng.standard_normal(n) * 0.02).astype(np.float32) for n in Ns] if epi == 0 else None
    res = rng.standard_normal((M, Ns[0])).astype(np.float32) if epi == 1 else None
    y = cu.gemv_fused([dev(v) for v in Ws], dev(x), norm_w=dev(w) if w is not None else None, eps=1e-6,
                      biases=[dev(b) for b in bs] if bs else None, epilogue=epi, residual=dev(res) if res is not None else None).cpu().numpy()
    xo = oracle.only.rmsnorm(x, w, 1e-6) if w is not None else x
    if epi == 0:
        want = np.concatenate([oracle.linear(Wi, xo, bi) for Wi, bi in zip(Ws, bs)], axis=1)
    elif epi == 1:
        want = oracle.linear(Ws[0], xo) + res
    else:
        want = oracle.silu(oracle.linear(Ws[0], xo)) * oracle.linear(Ws[1], xo)
    lin_check(y, want)
    # the same rows through mllmcu_linear (op-by-op path) take the same kernel: bit-identical for a single matrix
    if epi == 0 and len(Ns) == 1:
        xn = cu.rmsnorm(dev(x), dev(w), 1e-6)
        assert np.array_equal(bits(cu.linear(dev(Ws[0]), xn, dev(bs[0])).cpu().numpy()), bits(y))


def test_gemv_mma_exact_for_integer_data(cu):
    """Activations and weights chosen so every product is an integer and every scale a power of two: the fp32 result must be EXACT,
    which it is only if each mma.sync block sum is the exact integer Sum((nibble - 8) * q8)."""
    import torch
    rng = np.random.default_rng(3)
    M, N, K = 16, 64, 1024
    q = rng.integers(0, 16, (N, K // 32, 32)).astype(np.uint8)
    blk = np.empty((N, K // 32, 18), dtype=np.uint8)
    blk[..., 0:2] = np.array([1.0], dtype=np.float16).view(np.uint8)           # d_w = 1
    blk[..., 2:] = q[..., :16] | (q[..., 16:] << 4)
    xi = rng.integers(-127, 128, (M, K)).astype(np.float32)
    xi[:, ::32] = 127.0                                                         # amax = 127 in every block: d_x = 1, q8 = x
    y = cu.gemv_fused([dev(blk.reshape(N, -1))], dev(xi)).cpu().numpy()
    want = xi.astype(np.int64) @ (q.reshape(N, K).astype(np.int64) - 8).T
    assert np.array_equal(y.astype(np.int64), want) and np.abs(want).max() < 2 ** 24
