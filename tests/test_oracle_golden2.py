"""CPU tests: the oracle restatements of SURVEY 8a rows a4/a5/a7/a10/a11/a15/a17 (k-quants, the Q4_0_4_4 interleave,
eager softmax, GELU tables, M-RoPE / vision RoPE, patch-embed convolutions) against tests/golden/ref_vectors2.npz
(generated from the reference's own compiled sources by tests/golden/make_golden2.py) and, where oracle/_ref exists,
against the live compiled reference."""
import os

import numpy as np
import pytest

from oracle import pyoracle as po

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def bits(a):
    return np.ascontiguousarray(a, dtype=np.float32).view(np.uint32)


@pytest.fixture(scope="module")
def g2():
    return np.load(os.path.join(ROOT, "tests", "golden", "ref_vectors2.npz"))


def test_q8_K_blocks_bit_exact(oracle, g2):
    got = oracle.quantize_q8_K(g2["q8k_x"]).reshape(-1, po.Q8_K_BYTES)
    want = g2["q8k_blocks"].reshape(-1, po.Q8_K_BYTES)
    # an all-zero super-block leaves bsums unwritten in the reference (QuantizeQ8.cpp:229-234): compare d and qs there
    zero = np.all(want[:, :260] == 0, axis=1)
    assert np.array_equal(got[~zero], want[~zero])
    assert np.array_equal(got[zero][:, :260], want[zero][:, :260])


def test_q4_K_q6_K_dequant_bit_exact(oracle, g2):
    assert np.array_equal(bits(oracle.dequantize_kq(g2["q4k_blocks"], 512, po.Q4_K)), bits(g2["q4k_deq"]))
    assert np.array_equal(bits(oracle.dequantize_kq(g2["q6k_blocks"], 512, po.Q6_K)), bits(g2["q6k_deq"]))


def test_linear_k_quants_bit_exact(oracle, g2):
    """the lane-exact restatement of the AVX2 bodies reproduces the reference's fp32 results bit for bit"""
    assert np.array_equal(bits(oracle.linear_kq(g2["q4k_blocks"], g2["q8k_x"], g2["kq_bias"], po.Q4_K)), bits(g2["lin_q4k"]))
    assert np.array_equal(bits(oracle.linear_kq(g2["q6k_blocks"], g2["q8k_x"], None, po.Q6_K)), bits(g2["lin_q6k"]))


def test_k_quant_integer_sums_match_dequantised_dot(oracle, g2):
    xq = oracle.quantize_q8_K(g2["q8k_x"])
    for n in (0, 5, 23):
        for m in (0, 2, 4):
            sumi, summ = oracle.only.block_isums_q4_K_q8_K(g2["q4k_blocks"][n], xq[m], 512)
            blk = g2["q4k_blocks"][n].reshape(-1, 144)
            xb = xq[m].reshape(-1, 292)
            for i in range(2):
                sc = np.zeros(8, np.int64); mn = np.zeros(8, np.int64)
                q = blk[i, 4:16].astype(np.int64)
                for j in range(8):
                    if j < 4:
                        sc[j], mn[j] = q[j] & 63, q[j + 4] & 63
                    else:
                        sc[j] = (q[j + 4] & 0xF) | ((q[j - 4] >> 6) << 4)
                        mn[j] = (q[j + 4] >> 4) | ((q[j] >> 6) << 4)
                qs = blk[i, 16:].astype(np.int64).reshape(4, 32)
                w = np.stack([np.stack([qs[j] & 0xF, qs[j] >> 4]) for j in range(4)]).reshape(8, 32)
                q8 = xb[i, 4:260].view(np.int8).astype(np.int64).reshape(8, 32)
                assert sumi[i] == int((sc[:, None] * w * q8).sum())
                assert summ[i] == int((mn * q8.sum(axis=1)).sum())


def test_q4_0_4x4_interleave(oracle, g2):
    assert np.array_equal(oracle.only.q4_0_4x4_pack(g2["q44_rows"], 128), g2["q44_packed"])
    assert np.array_equal(oracle.only.q4_0_4x4_unpack(g2["q44_packed"], 8, 128), g2["q44_rows"])
    # the interleaved gemv computes the same Linear (different fp32 order)
    y = oracle.linear(g2["q44_rows"], g2["q44_x"][None, :])[0]
    assert np.abs(y - g2["q44_y"]).max() <= 2e-6 * np.abs(y).max()


def test_softmax_bit_exact(oracle, g2):
    assert np.array_equal(bits(oracle.softmax(g2["sm_x"])), bits(g2["sm_y"]))
    got = oracle.softmax(g2["sm_x"][:, :, :8].copy(), causal=True, old_dim=3)
    assert np.array_equal(bits(got), bits(g2["sm_y_causal"]))
    assert np.all(got[:, 0, 4:] == 0) and np.allclose(got.sum(-1), 1, atol=1e-6)


def test_gelu_tables_bit_exact(oracle, g2):
    for kind, tn, yn in ((0, "gelu_table", "gelu_y"), (1, "gelu_quick_table", "gelu_quick_y")):
        t = oracle.gelu_table(kind)
        assert np.array_equal(t, g2[tn])
        assert np.array_equal(bits(oracle.only.gelu(t, g2["gelu_x"])), bits(g2[yn]))


def test_conv2d_patch_embed(oracle, g2):
    for yn, st, b in (("conv_y", (14, 14), g2["conv_b"]), ("conv_y_s7", (7, 7), None)):
        y = oracle.conv2d_valid(g2["conv_x"], g2["conv_w"], b, st)
        assert y.shape == g2[yn].shape
        assert np.abs(y - g2[yn]).max() <= 1e-5 * np.abs(g2[yn]).max()  # vec_dot_fp32 sums in 8-lane order


def test_conv3d_equals_conv2d_over_flattened_time(oracle):
    rng = np.random.default_rng(3)
    x = rng.standard_normal((1, 3, 2, 28, 28)).astype(np.float32)
    w = (rng.standard_normal((8, 3, 2, 14, 14)) * 0.05).astype(np.float32)
    y3 = oracle.only.conv3d_valid(x, w, None, (2, 14, 14))
    y2 = oracle.conv2d_valid(x.reshape(1, 6, 28, 28), w.reshape(8, 6, 14, 14), None, (14, 14))
    assert np.array_equal(bits(y3.reshape(y2.shape)), bits(y2))
    ycl = oracle.only.conv3d_valid(x, w, None, (2, 14, 14), channels_last=True)
    assert np.array_equal(ycl, np.moveaxis(y3, 1, -1))


def test_mrope_text_positions_reduce_to_hf_rope(oracle):
    """with t = h = w position ids M-RoPE is plain HF RoPE (CPUMultimodalRoPE.cpp:62-81 picks the same column from every axis)"""
    S, D = 11, 128
    pos = np.tile(np.arange(5, 5 + S, dtype=np.float32), (3, 1))
    s, c = oracle.only.mrope_tables(pos, D, 1e6, [16, 24, 24, 16, 24, 24])
    sh, ch = oracle.only.rope_tables(4, 5 + S, D, 1e6)
    assert np.array_equal(bits(s), bits(sh[5:])) and np.array_equal(bits(c), bits(ch[5:]))
    # distinct axes: column blocks come from axes t,h,w,t,h,w
    pos = np.stack([np.arange(S), 2 * np.arange(S), 3 * np.arange(S)]).astype(np.float32)
    s, c = oracle.only.mrope_tables(pos, D, 1e6, [16, 24, 24, 16, 24, 24])
    theta = (1.0 / np.power(1e6, 2.0 * np.arange(D // 2) / D)).astype(np.float32)
    axis = np.repeat([0, 1, 2, 0, 1, 2], [16, 24, 24, 16, 24, 24])
    want = np.sin((theta[np.arange(D) % 64][None, :] * pos[axis].T).astype(np.float32))
    assert np.allclose(s, want, atol=1e-6)


def test_vision_rope_rotation(oracle):
    rng = np.random.default_rng(9)
    S, H, D = 6, 2, 80
    x = rng.standard_normal((S, H, D)).astype(np.float32)
    ang = (rng.uniform(0, 60, (S, D // 2))).astype(np.float32)
    y = oracle.only.rope_vision(x, ang)
    sn, cs = np.sin(ang)[:, None, :], np.cos(ang)[:, None, :]
    a, b = x[..., : D // 2], x[..., D // 2:]
    assert np.allclose(y[..., : D // 2], a * cs - b * sn, atol=2e-6) and np.allclose(y[..., D // 2:], a * sn + b * cs, atol=2e-6)


def test_live_reference_agrees_when_built(oracle, ref_lib):
    if ref_lib is None:
        pytest.skip("oracle/_ref not built here")
    rng = np.random.default_rng(123)
    x = (rng.standard_normal((5, 1024)) * rng.uniform(0.01, 5, (5, 1))).astype(np.float32)
    assert np.array_equal(oracle.quantize_q8_K(x), ref_lib.quantize_q8_K(x))
    Wk = rng.integers(0, 256, (9, 1024 // 256 * 144), dtype=np.uint8)  # any bytes are a valid Q4_K row once d/dmin are finite
    Wk = Wk.reshape(9, -1, 144)
    Wk[:, :, 0:4] = np.array([0.01, 0.005], dtype=np.float16).view(np.uint8)
    Wk = Wk.reshape(9, -1)
    assert np.array_equal(bits(oracle.linear_kq(Wk, x)), bits(ref_lib.linear_kq(Wk, x)))
    W6 = rng.integers(0, 256, (9, 1024 // 256 * 210), dtype=np.uint8).reshape(9, -1, 210)
    W6[:, :, 208:210] = np.array([0.01], dtype=np.float16).view(np.uint8)
    W6 = W6.reshape(9, -1)
    assert np.array_equal(bits(oracle.linear_kq(W6, x, wtype=po.Q6_K)), bits(ref_lib.linear_kq(W6, x, wtype=po.Q6_K)))
    sx = (rng.standard_normal((3, 7, 29)) * 4).astype(np.float32)
    assert np.array_equal(bits(oracle.softmax(sx)), bits(ref_lib.softmax(sx)))
    assert np.array_equal(bits(oracle.softmax(sx[:, :, :9].copy(), True, 2)), bits(ref_lib.softmax(sx[:, :, :9].copy(), True, 2)))
