"""ncu / timing driver: prefill attention at the LLaMA-2-7B shape (argv: S [H D])."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from mllm_b200 import capi  # noqa: E402

S = int(sys.argv[1]) if len(sys.argv) > 1 else 512
H = int(sys.argv[2]) if len(sys.argv) > 2 else 32
D = int(sys.argv[3]) if len(sys.argv) > 3 else 128
q = torch.randn((1, S, H, D), device="cuda")
k = torch.randn((1, S, H, D), device="cuda", dtype=torch.float16)
v = torch.randn((1, S, H, D), device="cuda", dtype=torch.float16)
for _ in range(4):
    capi.flash_attn_f16kv(q, k, v)
torch.cuda.synchronize()
print("ok")
