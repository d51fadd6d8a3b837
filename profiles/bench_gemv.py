"""Micro-benchmark of the decode Linear kernels (not a bench.py number): GB/s per shape, cold weights."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from mllm_b200 import capi  # noqa: E402

SHAPES = [("floor 296x4096 (1 row pair per SM)", [296], 4096, 0), ("o/q 4096x4096", [4096], 4096, 0), ("qkv 3x4096x4096", [4096] * 3, 4096, 0), ("gate+up 2x11008x4096", [11008] * 2, 4096, 2),
          ("down 4096x11008", [4096], 11008, 1), ("lm_head 32000x4096", [32000], 4096, 0)]


def run(tag, Ns, K, epi, M=1, fused=True, pdl=False):
    wb = sum(Ns) * (K // 32) * 18
    nbuf = max(2, int(200e6 // wb) + 1)
    bufs = [[torch.randint(0, 255, (n, K // 32 * 18), dtype=torch.uint8, device="cuda") for n in Ns] for _ in range(nbuf)]
    x = torch.randn((M, K), device="cuda")
    nw = torch.ones(K, device="cuda")
    res = torch.randn((M, Ns[0]), device="cuda")

    def call(ws):
        if fused:
            capi.gemv_fused(ws, x, norm_w=nw if epi != 1 else None, epilogue=epi, residual=res if epi == 1 else None, pdl=pdl)
        else:
            for w in ws:
                capi.linear(w, x)
    for i in range(3):
        call(bufs[i % nbuf])
    reps = 4 * nbuf
    # the launches are captured into one CUDA graph so the timing is device time, not Python/ctypes dispatch
    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=side):
            for i in range(reps):
                call(bufs[i % nbuf])
        g.replay()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        e0.record()
        g.replay()
        e1.record()
        torch.cuda.synchronize()
    t = e0.elapsed_time(e1) / reps * 1e-3
    print(f"{tag:28s} M={M} fused={int(fused)} pdl={int(pdl)} rpw={os.environ.get('MLLMCU_GEMV_RPW', 'auto'):>4s}  {t * 1e6:8.2f} us  {wb / t / 1e9:8.1f} GB/s", flush=True)


if __name__ == "__main__":
    mode = sys.argv[1] if len(sys.argv) > 1 else "all"
    for tag, Ns, K, epi in SHAPES:
        if mode == "all":
            run(tag, Ns, K, epi, fused=False)
        run(tag, Ns, K, epi, fused=True)
        run(tag, Ns, K, epi, fused=True, pdl=True)
    if mode == "all":
        for M in (2, 4, 8):
            run("gate+up 2x11008x4096", [11008] * 2, 4096, 2, M=M)
