"""CPU tests: the oracle restatement (oracle/oracle.c) against the golden vectors that
tests/golden/make_golden.py generated from the reference's own compiled sources, and
against the live compiled reference (oracle/_ref) where that exists."""
import numpy as np
import pytest

from oracle import pyoracle as po


def bits(a):
    return np.ascontiguousarray(a, dtype=np.float32).view(np.uint32)


def test_q8_0_blocks_bit_exact(oracle, golden):
    assert np.array_equal(oracle.quantize_q8_0(golden["q8_x"]), golden["q8_blocks"])


def test_q4_0_quant_dequant_bit_exact(oracle, golden):
    blocks = oracle.quantize_q4_0(golden["q4_w"])
    assert np.array_equal(blocks, golden["q4_blocks"])
    assert np.array_equal(bits(oracle.dequantize_q4_0(blocks, 256)), bits(golden["q4_deq"]))


def test_linear_q4_0_bit_exact(oracle, golden):
    W, x, b = golden["q4_blocks"], golden["q8_x"], golden["lin_bias"]
    assert np.array_equal(bits(oracle.linear(W, x)), bits(golden["lin_y"]))
    assert np.array_equal(bits(oracle.linear(W, x, b)), bits(golden["lin_y_bias"]))
    assert np.array_equal(oracle.linear(W, x, b, out_fp16=True), golden["lin_y_bias_f16"])


def test_linear_q8_0_bit_exact(oracle, golden):
    assert np.array_equal(bits(oracle.linear(golden["q8w_blocks"], golden["q8_x"], wtype=po.Q8_0)), bits(golden["lin8_y"]))


def test_silu_bit_exact(oracle, golden):
    assert np.array_equal(bits(oracle.silu(golden["silu_x"])), bits(golden["silu_y"]))


def test_fp16_rne_bit_exact(oracle, golden):
    assert np.array_equal(oracle.only.fp32_to_fp16_bits(golden["f16_x"]), golden["f16_bits"])
    # and against numpy's IEEE conversion
    assert np.array_equal(golden["f16_bits"], golden["f16_x"].astype(np.float16).view(np.uint16))


@pytest.mark.parametrize("tag", ["dec", "pre", "app"])
def test_flash_attention_tolerance(oracle, golden, tag):
    o = oracle.flash_attn_f16kv(golden[f"fa_{tag}_q"], golden[f"fa_{tag}_k"], golden[f"fa_{tag}_v"])
    ref = golden[f"fa_{tag}_o"]
    # accumulation order differs from the AVX tiles: 2e-6 absolute on O(1) outputs
    assert np.abs(o - ref).max() <= 2e-6 * max(1.0, np.abs(ref).max())


def test_block_isums_match_lane_sums(oracle, golden):
    """sum over the 8 AVX lanes == dequantised integer dot, per block."""
    W, xq = golden["q4_blocks"], golden["q8_blocks"]
    K = 256
    for n in (0, 7, 39):
        for m in (0, 2, 5):
            isum = oracle.only.block_isums_q4_0_q8_0(W[n], xq[m], K)
            wq = W[n].reshape(-1, 18)[:, 2:]
            lo = (wq & 0x0F).astype(np.int32) - 8
            hi = (wq >> 4).astype(np.int32) - 8
            wi = np.concatenate([lo, hi], axis=1)
            xi = xq[m].reshape(-1, 34)[:, 2:].view(np.int8).astype(np.int32)
            assert np.array_equal(isum, (wi * xi).sum(axis=1))


def test_rmsnorm_against_numpy(oracle):
    rng = np.random.default_rng(5)
    x = rng.standard_normal((7, 2048)).astype(np.float32)
    w = (1 + 0.02 * rng.standard_normal(2048)).astype(np.float32)
    y = oracle.only.rmsnorm(x, w, 1e-6)
    rms = (1.0 / np.sqrt((x.astype(np.float64) ** 2).mean(axis=1).astype(np.float32) + np.float32(1e-6))).astype(np.float32)
    want = (x * rms[:, None]).astype(np.float32) * w
    assert np.array_equal(bits(y), bits(want))


@pytest.mark.parametrize("rtype", [po_t for po_t in (2, 4)])
def test_rope_tables_and_rotation(oracle, rtype):
    rng = np.random.default_rng(6)
    D, pos_max = 64, 40
    s, c = oracle.only.rope_tables(rtype, pos_max, D, 10000.0)
    theta = (1.0 / np.power(10000.0, 2.0 * np.arange(D // 2) / D)).astype(np.float32)
    t = (np.arange(pos_max, dtype=np.float32)[:, None] * theta[None, :]).astype(np.float32)
    if rtype == 2:
        assert np.allclose(s[:, 0::2], np.sin(t), atol=1e-6) and np.array_equal(s[:, 0::2], s[:, 1::2])
    else:
        assert np.allclose(s[:, : D // 2], np.sin(t), atol=1e-6) and np.array_equal(s[:, : D // 2], s[:, D // 2:])
    x = rng.standard_normal((5, 3, D)).astype(np.float32)
    y = oracle.only.rope_apply(rtype, x, 7, s, c)
    # rotation preserves pair norms
    if rtype == 2:
        n_in = x[..., 0::2] ** 2 + x[..., 1::2] ** 2
        n_out = y[..., 0::2] ** 2 + y[..., 1::2] ** 2
    else:
        n_in = x[..., : D // 2] ** 2 + x[..., D // 2:] ** 2
        n_out = y[..., : D // 2] ** 2 + y[..., D // 2:] ** 2
    assert np.allclose(n_in, n_out, rtol=1e-5, atol=1e-6)
    y16 = oracle.only.rope_apply(rtype, x, 7, s, c, out_fp16=True)
    assert np.array_equal(y16, y.astype(np.float16).view(np.uint16))


def test_live_reference_agrees_when_built(oracle, ref_lib):
    """Where oracle/_ref exists (build container, or shipped to the GPU box) compare live on fresh inputs."""
    if ref_lib is None:
        pytest.skip("oracle/_ref not built here")
    rng = np.random.default_rng(77)
    x = (rng.standard_normal((9, 512)) * rng.uniform(0.01, 5, (9, 1))).astype(np.float32)
    w = (rng.standard_normal((33, 512)) * 0.02).astype(np.float32)
    assert np.array_equal(oracle.quantize_q8_0(x), ref_lib.quantize_q8_0(x))
    W = oracle.quantize_q4_0(w)
    assert np.array_equal(W, ref_lib.quantize_q4_0(w))
    assert np.array_equal(bits(oracle.linear(W, x)), bits(ref_lib.linear(W, x)))
    z = (rng.standard_normal(4096) * 5).astype(np.float32)
    assert np.array_equal(bits(oracle.silu(z)), bits(ref_lib.silu(z)))
