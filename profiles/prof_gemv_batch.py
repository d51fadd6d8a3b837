"""ncu / timing driver: the fused decode Linear at batch M on Qwen1.5-1.8B shapes (argv: M [shape])."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from mllm_b200 import capi  # noqa: E402

M = int(sys.argv[1]) if len(sys.argv) > 1 else 8
which = sys.argv[2] if len(sys.argv) > 2 else "all"
SHAPES = {"qkv": ([2048] * 3, 2048, 0), "o": ([2048], 2048, 1), "gateup": ([5504] * 2, 2048, 2), "down": ([2048], 5504, 1), "lm_head": ([151936], 2048, 0)}
for name, (Ns, K, epi) in SHAPES.items():
    if which not in ("all", name):
        continue
    nbuf = 4
    bufs = [[torch.randint(0, 255, (n, K // 32 * 18), dtype=torch.uint8, device="cuda") for n in Ns] for _ in range(nbuf)]
    for b in bufs:
        for w in b:
            w.view(w.shape[0], K // 32, 18)[:, :, 1] = 0x20
    x = torch.randn((M, K), device="cuda")
    nw = torch.ones(K, device="cuda")
    res = torch.randn((M, Ns[0]), device="cuda")

    def call(i):
        capi.gemv_fused(bufs[i % nbuf], x, norm_w=nw if epi != 1 else None, epilogue=epi, residual=res if epi == 1 else None)
    for i in range(3):
        call(i)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    for i in range(20):
        call(i)
    e1.record()
    torch.cuda.synchronize()
    t = e0.elapsed_time(e1) / 20 * 1e-3
    wb = sum(Ns) * (K // 32) * 18
    print(f"M={M} {name:8s} N={sum(Ns) if epi != 2 else Ns[0]:7d} K={K:5d}: {t * 1e6:8.1f} us  {wb / t / 1e9:7.1f} GB/s  {2.0 * M * sum(Ns) * K / t / 1e12:6.2f} Top/s", flush=True)
