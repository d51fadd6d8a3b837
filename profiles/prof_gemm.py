"""ncu / timing driver for the prefill Linear (tcgen05 path): argv M N K [reps]."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from mllm_b200 import capi  # noqa: E402

M, N, K = (int(a) for a in sys.argv[1:4]) if len(sys.argv) > 3 else (512, 4096, 4096)
reps = int(sys.argv[4]) if len(sys.argv) > 4 else 5
W = torch.randint(0, 255, (N, K // 32 * 18), dtype=torch.uint8, device="cuda")
Wv = W.view(N, K // 32, 18)
Wv[:, :, 1] = 0x20  # sane fp16 scales
x = torch.randn((M, K), device="cuda")
for _ in range(3):
    capi.linear(W, x)
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
torch.cuda.synchronize()
e0.record()
for _ in range(reps):
    capi.linear(W, x)
e1.record()
torch.cuda.synchronize()
t = e0.elapsed_time(e1) / reps * 1e-3
print(f"M={M} N={N} K={K}: {t * 1e6:.1f} us per Linear (quantise + GEMM), {2.0 * M * N * K / t / 1e12:.1f} Tops")
