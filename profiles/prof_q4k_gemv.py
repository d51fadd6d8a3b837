"""One shape of the Q4_K decode GEMV for an ncu capture: gate|up of LLaMA-2-7B / LLaVA (2 x 11008 rows x K 4096, M = 1), 304-byte activation blocks.
usage: PYTHONPATH=. python profiles/prof_q4k_gemv.py"""
import torch
from mllm_b200 import capi as cu

dev = torch.device("cuda:0")
K, N = 4096, 11008
Ws = []
for _ in range(2):
    W = torch.randint(0, 256, (N, K // 256, 144), dtype=torch.uint8, device=dev)
    d = torch.full((N, K // 256, 1), 0.01, dtype=torch.float16, device=dev).view(torch.uint8)
    W[:, :, 0:2] = d
    W[:, :, 2:4] = d
    Ws.append(W.reshape(N, -1))
x = torch.randn((1, K), device=dev)
w = torch.ones((K,), device=dev)
for _ in range(5):
    xq = cu.prep_q8_K(x, 1, w, 1e-5, pdl=True, aligned=True)
    g, u = cu.gemv_q4k(Ws, xq, K, pdl=True, aligned=True)
    gq = cu.prep_q8_K(g, 2, u, pdl=True, aligned=True)
torch.cuda.synchronize()
print("ok", float(g.abs().mean()))
