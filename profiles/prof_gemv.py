"""ncu driver: a few launches of the fused decode GEMV for one shape (argv: down|gateup|o)."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from mllm_b200 import capi  # noqa: E402

which = sys.argv[1] if len(sys.argv) > 1 else "down"
SHAPES = {"qkv": ([4096] * 3, 4096, 0), "o": ([4096], 4096, 1), "gateup": ([11008] * 2, 4096, 2), "down": ([4096], 11008, 1), "lm_head": ([32000], 4096, 0)}
if which == "all":  # two launches per decode-step shape on distinct weight buffers (the second one is the one to read)
    for name, (Ns, K, epi) in SHAPES.items():
        x = torch.randn((1, K), device="cuda")
        nw = torch.ones(K, device="cuda")
        res = torch.randn((1, Ns[0]), device="cuda")
        for i in range(2):
            W = [torch.randint(0, 255, (n, K // 32 * 18), dtype=torch.uint8, device="cuda") for n in Ns]
            capi.gemv_fused(W, x, norm_w=nw if epi != 1 else None, epilogue=epi, residual=res if epi == 1 else None)
            torch.cuda.synchronize()
    print("ok")
    sys.exit(0)
Ns, K, epi = SHAPES[which]
bufs = [[torch.randint(0, 255, (n, K // 32 * 18), dtype=torch.uint8, device="cuda") for n in Ns] for _ in range(6)]
x = torch.randn((1, K), device="cuda")
nw = torch.ones(K, device="cuda")
res = torch.randn((1, Ns[0]), device="cuda")
for i in range(6):
    capi.gemv_fused(bufs[i], x, norm_w=nw if epi != 1 else None, epilogue=epi, residual=res if epi == 1 else None)
torch.cuda.synchronize()
print("ok")
