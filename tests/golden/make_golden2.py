"""Generate tests/golden/ref_vectors2.npz (rows a4/a5/a7/a15/a17 of SURVEY 8a) from the REFERENCE's own
compiled sources (oracle/_ref/libmllm_ref.so: ggml/{QuantizeQ8,QuantizeQ4,QuantizeQ6,VecDotQ4,VecDotQ6,GemmPack,
VecDotFP32}.cpp, ggml/Quantize.hpp, compute/ActivationFunction.cpp).

    make -C oracle && python tests/golden/make_golden2.py
"""
import ctypes as C
import os
import sys

import numpy as np

sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", ".."))
from oracle import pyoracle as po  # noqa: E402
from oracle.pyoracle import _p  # noqa: E402


def rows_kq(R, fn, w, bs):
    out = np.empty((w.shape[0], w.shape[1] // 256 * bs), dtype=np.uint8)
    for r in range(w.shape[0]):
        R.fn(fn)(_p(w[r]), _p(out[r]), C.c_int(w.shape[1]))
    return out


def main():
    R = po.ref()
    assert R is not None, "build oracle/_ref first (make -C oracle)"
    rng = np.random.default_rng(4321)
    out = {}
    # --- Q8_K activation rows: sign of the max element, ties on |x|, exact .5 products, an all-zero super-block
    x = (rng.standard_normal((6, 512)) * rng.uniform(0.01, 8, (6, 1))).astype(np.float32)
    x[1, :256] = 0.0
    x[2, :256] = np.arange(256, dtype=np.float32) * 0.25 - 32.0  # max = -32 -> iscale = 4 -> every product is an integer
    x[3, 5] = 9.5
    x[3, 9] = -9.5  # equal |x|: the first one decides the sign of iscale
    x[4, 256 + 7] = -x[4, 256:].max() * 3
    out["q8k_x"] = x
    out["q8k_blocks"] = R.quantize_q8_K(x)
    # --- Q4_K / Q6_K weights from the reference quantiser + their dots against the rows above
    w = (rng.standard_normal((24, 512)) * 0.02).astype(np.float32)
    w[0, :256] = 0.0
    out["kq_w"] = w
    out["q4k_blocks"] = rows_kq(R, "quantize_row_q4_K", w, po.Q4_K_BYTES)
    out["q6k_blocks"] = rows_kq(R, "quantize_row_q6_K", w, po.Q6_K_BYTES)
    out["q4k_deq"] = R.dequantize_kq(out["q4k_blocks"], 512, po.Q4_K)
    out["q6k_deq"] = R.dequantize_kq(out["q6k_blocks"], 512, po.Q6_K)
    bias = (rng.standard_normal(24) * 0.02).astype(np.float32)
    out["kq_bias"] = bias
    out["lin_q4k"] = R.linear_kq(out["q4k_blocks"], x, bias, po.Q4_K)
    out["lin_q6k"] = R.linear_kq(out["q6k_blocks"], x, None, po.Q6_K)
    # --- Q4_0_4_4: the 4-row interleave of the same rows the plain quantiser produces
    w4 = (rng.standard_normal((8, 128)) * 0.02).astype(np.float32)
    out["q44_w"] = w4
    packed = np.empty((8, 128 // 32 * 18), dtype=np.uint8)
    R.fn("quantize_q4_0_4x4")(_p(w4), _p(packed), 8, 128)
    out["q44_packed"] = packed
    out["q44_rows"] = R.quantize_q4_0(w4)
    xv = (rng.standard_normal(128) * 2).astype(np.float32)
    yv = np.empty(8, dtype=np.float32)
    R.fn("gemv_q4_0_4x4")(_p(packed), _p(xv), _p(yv), 8, 128)
    out["q44_x"], out["q44_y"] = xv, yv
    # --- softmax (eager attention): plain and causal with a past of 3, lengths that exercise the n%8 tail
    sx = (rng.standard_normal((2, 5, 21)) * 3).astype(np.float32)
    out["sm_x"] = sx
    out["sm_y"] = R.softmax(sx)
    out["sm_y_causal"] = R.softmax(sx[:, :, :8].copy(), causal=True, old_dim=3)
    # --- GELU tables + lookups
    out["gelu_table"] = R.gelu_table(0)
    out["gelu_quick_table"] = R.gelu_table(1)
    gx = (rng.standard_normal(300) * 3).astype(np.float32)
    gx[:6] = [0, -0.0, 70000.0, -70000.0, 1e-9, 6.1e-5]
    out["gelu_x"] = gx
    for kind, nm in ((0, "gelu_y"), (1, "gelu_quick_y")):
        y = np.empty_like(gx)
        R.fn("gelu")(kind, _p(gx), _p(y), gx.size)
        out[nm] = y
    # --- patch-embed convolution (stride = kernel) and an overlapping one
    cx = rng.standard_normal((1, 3, 28, 28)).astype(np.float32)
    cw = (rng.standard_normal((16, 3, 14, 14)) * 0.05).astype(np.float32)
    cb = rng.standard_normal(16).astype(np.float32)
    out["conv_x"], out["conv_w"], out["conv_b"] = cx, cw, cb
    out["conv_y"] = R.conv2d_valid(cx, cw, cb, (14, 14))
    out["conv_y_s7"] = R.conv2d_valid(cx, cw, None, (7, 7))
    path = os.path.join(os.path.dirname(__file__), "ref_vectors2.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
