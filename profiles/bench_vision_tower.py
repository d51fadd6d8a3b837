#!/usr/bin/env python
"""Device timing of a vision tower composed from the C-ABI kernels (no parity claim beyond tests/test_backend_ops_gpu.py's ViT block and
the per-kernel tests): BASELINE configs[2] CLIP ViT-L/14-336 as LLaVA-1.5 uses it (models/llava/modeling_llava.hpp:37-105: patch conv,
CLS + position embedding, pre-LayerNorm, 23 ViT blocks with QuickGELU, drop CLS, projector Linear-GELU-Linear into the LLM width),
Linears in Q4_K as the shipped LLaVA file quantises them, random-init weights, synthetic 3x336x336 image.
    python profiles/bench_vision_tower.py [--layers 23]"""
import argparse
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from mllm_b200 import capi, mllm_file as mf  # noqa: E402

Q4_K = 12


def q4k(rows, cols, rng):
    return torch.from_numpy(mf.quantize_q4_K_np((0.02 * rng.standard_normal((rows, cols))).astype(np.float32))).cuda()


def vec(n, rng, base=0.0):
    return torch.from_numpy((base + 0.02 * rng.standard_normal(n)).astype(np.float32)).cuda()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--layers", type=int, default=23)
    ap.add_argument("--reps", type=int, default=5)
    a = ap.parse_args()
    rng = np.random.default_rng(7)
    Hd, H, F, P, IMG, LLM = 1024, 16, 4096, 14, 336, 4096
    D, S = Hd // H, (IMG // P) ** 2 + 1
    conv_w = torch.from_numpy((0.02 * rng.standard_normal((Hd, 3, P, P))).astype(np.float32)).cuda()
    cls, pos = vec(Hd, rng), torch.from_numpy((0.02 * rng.standard_normal((S, Hd))).astype(np.float32)).cuda()
    pre = (vec(Hd, rng, 1.0), vec(Hd, rng))
    blocks = []
    for _ in range(a.layers):
        blocks.append(dict(ln1=(vec(Hd, rng, 1.0), vec(Hd, rng)), ln2=(vec(Hd, rng, 1.0), vec(Hd, rng)),
                           q=(q4k(Hd, Hd, rng), vec(Hd, rng)), k=(q4k(Hd, Hd, rng), vec(Hd, rng)), v=(q4k(Hd, Hd, rng), vec(Hd, rng)),
                           o=(q4k(Hd, Hd, rng), vec(Hd, rng)), fc1=(q4k(F, Hd, rng), vec(F, rng)), fc2=(q4k(Hd, F, rng), vec(Hd, rng))))
    proj1, proj2 = (q4k(LLM, Hd, rng), vec(LLM, rng)), (q4k(LLM, LLM, rng), vec(LLM, rng))
    # fp16 lookup tables over all 65 536 halves, as init_table_gelu_f16 / gelu_quick build them (ggml/Quantize.hpp:74-133); numpy is enough here
    hv = np.arange(65536, dtype=np.uint16).view(np.float16).astype(np.float32)
    with np.errstate(all="ignore"):
        tq = (hv / (1.0 + np.exp(-1.702 * hv))).astype(np.float16).view(np.uint16)
        tg = (0.5 * hv * (1.0 + np.tanh(0.79788456080286535587989211986876 * hv * (1.0 + 0.044715 * hv * hv)))).astype(np.float16).view(np.uint16)
    quick, gelu = torch.from_numpy(tq.copy()).cuda(), torch.from_numpy(tg.copy()).cuda()
    img = torch.rand((1, 3, IMG, IMG), device="cuda")

    def lin(w, x):
        return capi.linear(w[0], x, w[1], wtype=Q4_K)

    def forward():
        e = capi.conv2d_valid(img, conv_w, None, (P, P), channels_last=True)[0]  # [576, 1024]: patches x channels (transpose + flatten folded)
        x = torch.cat([cls[None], e], 0) + pos                                  # cat + position embedding
        x = capi.layernorm(x, pre[0], pre[1], 1e-6)
        for b in blocks:
            h = capi.layernorm(x, b["ln1"][0], b["ln1"][1], 1e-5)
            q, k, v = (lin(b[n], h).view(1, S, H, D) for n in ("q", "k", "v"))
            at = capi.flash_attn_f32kv(q, k, v, causal=False).view(S, Hd)
            x = capi.add(lin(b["o"], at), x)
            h = capi.layernorm(x, b["ln2"][0], b["ln2"][1], 1e-5)
            x = capi.add(lin(b["fc2"], capi.gelu_table(lin(b["fc1"], h), quick)), x)
        f = x[1:].contiguous()                                                  # clip({}, {}, {1, clip_len}, {})
        return lin(proj2, capi.gelu_table(lin(proj1, f), gelu))

    out = forward()
    assert out.shape == (S - 1, LLM) and torch.isfinite(out).all()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    best = 1e9
    for _ in range(a.reps):
        torch.cuda.synchronize()
        e0.record()
        forward()
        e1.record()
        torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    flops = a.layers * (2 * S * (4 * Hd * Hd + 2 * Hd * F) + 4 * S * S * Hd) + 2 * (S - 1) * (Hd * LLM + LLM * LLM)
    print(f"CLIP ViT-L/14-336 tower ({a.layers} blocks, {S} tokens, Q4_K Linears) + projector: {best:.2f} ms per image, {flops / best / 1e9:.1f} Tflop/s-equivalent")


if __name__ == "__main__":
    main()
