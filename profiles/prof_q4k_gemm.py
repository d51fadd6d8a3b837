"""Timing of the Q4_K Linear at prefill shapes (LLaVA-1.5-7B: 608 rows; CLIP tower: 577 rows, K 1024 / 4096): argv = [M]."""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from mllm_b200 import capi  # noqa: E402

M = int(sys.argv[1]) if len(sys.argv) > 1 else 608
for N, K in ((4096, 4096), (11008, 4096), (4096, 11008), (1024, 1024), (4096, 1024), (1024, 4096)):
    W = torch.randint(0, 255, (N, K // 256 * 144), dtype=torch.uint8, device="cuda")
    Wv = W.view(N, K // 256, 144)
    Wv[:, :, 0:4] = torch.tensor([0x00, 0x10, 0x00, 0x14], dtype=torch.uint8, device="cuda")  # d, dmin: small finite fp16
    x = torch.randn((M, K), device="cuda")
    for _ in range(3):
        y = capi.linear(W, x, wtype=capi.TYPE_Q4_K)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(10):
        y = capi.linear(W, x, wtype=capi.TYPE_Q4_K)
    e1.record()
    torch.cuda.synchronize()
    t = e0.elapsed_time(e1) / 10 * 1e-3
    print(f"M={M} N={N:6d} K={K:6d}: {t * 1e6:9.1f} us per Linear (quantise + GEMM)  {2.0 * M * N * K / t / 1e12:7.2f} Tops", flush=True)
