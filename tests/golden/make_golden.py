"""Generate tests/golden/ref_vectors.npz from the REFERENCE's own compiled sources.

Run in the build container (where /root/reference exists):
    make -C oracle && python tests/golden/make_golden.py
The vectors are outputs of oracle/_ref/libmllm_ref.so, i.e. of
/root/reference/mllm/backends/cpu/third_party/ggml/{QuantizeQ8,QuantizeQ4,VecDotQ4,VecDotQ8}.cpp,
compute/ActivationFunction.cpp and compute/FlashAttention2.hpp compiled with
-mavx2 -mf16c -mfma (oracle/Makefile).  They pin oracle.c wherever the GPU box
has no /root/reference.
"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", ".."))
from oracle import pyoracle as po  # noqa: E402


def main():
    R = po.ref()
    assert R is not None, "build oracle/_ref first (make -C oracle)"
    rng = np.random.default_rng(1234)
    out = {}
    # --- Q8_0 activation rows: scales, zeros, exact .5 ties, large/small magnitudes
    x = (rng.standard_normal((12, 256)) * rng.uniform(0.01, 8, (12, 1))).astype(np.float32)
    x[1, :32] = 0.0
    x[2, :32] = np.arange(32, dtype=np.float32) * 0.5
    x[2, 31] = 63.5  # amax 63.5 -> id = 2 -> every element lands on .0/.5 -> ties-to-even
    x[3, 32:64] = -x[3, 32:64].max()
    x[4] *= 1e-6
    x[5] *= 3e4
    out["q8_x"] = x
    out["q8_blocks"] = R.quantize_q8_0(x)
    # --- Q4_0 weights
    w = (rng.standard_normal((40, 256)) * 0.02).astype(np.float32)
    w[0, :32] = 0.0
    out["q4_w"] = w
    out["q4_blocks"] = R.quantize_q4_0(w)
    out["q4_deq"] = R.dequantize_q4_0(out["q4_blocks"], 256)
    # --- Linear (Q4_0 weights): fp32 out, bias, fp16 out
    bias = (rng.standard_normal(40) * 0.02).astype(np.float32)
    out["lin_bias"] = bias
    out["lin_y"] = R.linear(out["q4_blocks"], x)
    out["lin_y_bias"] = R.linear(out["q4_blocks"], x, bias)
    out["lin_y_bias_f16"] = R.linear(out["q4_blocks"], x, bias, out_fp16=True)
    # --- Linear (Q8_0 weights)
    w8 = R.quantize_q8_0(w)
    out["q8w_blocks"] = w8
    out["lin8_y"] = R.linear(w8, x, wtype=po.Q8_0)
    # --- SiLU incl. the polynomial exp's special ranges
    z = (rng.standard_normal(256) * 4).astype(np.float32)
    z[:12] = [0, -0.0, 100, -100, 88.5, -88.5, -104, 1e-30, -200, 200, 127.5, -127.5]
    out["silu_x"] = z
    out["silu_y"] = R.silu(z)
    # --- fp16 conversion
    f = np.concatenate([(rng.standard_normal(64) * s).astype(np.float32) for s in (1e-8, 1e-5, 1e-3, 1, 100, 70000)])
    f = np.concatenate([f, np.array([0, -0.0, 65504, 65519.99, 65520, 1e-7, 5.96e-8, 2.98e-8, 2.9802322e-8, 6.1e-5], dtype=np.float32)])
    out["f16_x"] = f
    out["f16_bits"] = np.array([R.lib.ref_fp32_to_fp16(float(v)) for v in f], dtype=np.uint16)
    # --- flash attention, fp16 KV (prefill Sk%4==0 because of the reference's Tc_left defect,
    #     FlashAttention2.hpp:1274 `seq_size_k % Tc`; decode any Sk)
    P = po.port()
    for tag, (Sq, Sk, Hq, Hkv, D) in dict(dec=(1, 37, 8, 2, 64), pre=(16, 16, 4, 4, 128), app=(4, 12, 4, 2, 64)).items():
        q = rng.standard_normal((Sq, Hq, D)).astype(np.float32)
        k = P.only.fp32_to_fp16_bits(rng.standard_normal((Sk, Hkv, D)).astype(np.float32))
        v = P.only.fp32_to_fp16_bits(rng.standard_normal((Sk, Hkv, D)).astype(np.float32))
        out[f"fa_{tag}_q"], out[f"fa_{tag}_k"], out[f"fa_{tag}_v"] = q, k, v
        out[f"fa_{tag}_o"] = R.flash_attn_f16kv(q, k, v, threads=1)
    path = os.path.join(os.path.dirname(__file__), "ref_vectors.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
