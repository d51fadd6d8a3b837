"""Device timing of the prefill kernels on LLaMA-2-7B shapes: both tcgen05 Linear kernels (scaled-operand default vs
MLLMCU_TC_EXACT=1) and both attention kernels (tensor-core default vs MLLM_CUDA_ATTN_TC=0).  CUDA events on torch's current
stream (the C ABI is called with it), weights rotated over > L2-sized buffers."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from mllm_b200 import capi  # noqa: E402


def timed(fn, reps=10):
    for _ in range(3):
        fn(0)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    for i in range(reps):
        fn(i)
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps * 1e-3


M = int(sys.argv[1]) if len(sys.argv) > 1 else 512
for N, K in [(4096, 4096), (11008, 4096), (4096, 11008), (12288, 4096)]:
    nbuf = max(2, int(200e6 // (N * K // 32 * 18)) + 1)
    Ws = []
    for _ in range(nbuf):
        W = torch.randint(0, 255, (N, K // 32 * 18), dtype=torch.uint8, device="cuda")
        W.view(N, K // 32, 18)[:, :, 1] = 0x20
        Ws.append(W)
    x = torch.randn((M, K), device="cuda")
    for mode in ("0", "1"):
        os.environ["MLLMCU_TC_EXACT"] = mode
        t = timed(lambda i: capi.linear(Ws[i % nbuf], x))
        print(f"linear M={M} N={N} K={K} exact={mode}: {t * 1e6:8.1f} us  {2.0 * M * N * K / t / 1e12:7.1f} Top/s-equivalent", flush=True)
os.environ.pop("MLLMCU_TC_EXACT", None)

for S, H, D in [(512, 32, 128), (2048, 32, 128), (512, 32, 64)]:
    q = torch.randn((1, S, H, D), device="cuda")
    k = torch.randn((1, S, H, D), device="cuda", dtype=torch.float16)
    v = torch.randn((1, S, H, D), device="cuda", dtype=torch.float16)
    for mode in ("1", "0"):
        os.environ["MLLM_CUDA_ATTN_TC"] = mode
        t = timed(lambda i: capi.flash_attn_f16kv(q, k, v))
        print(f"attention S={S} H={H} D={D} tc={mode}: {t * 1e6:8.1f} us  {4.0 * S * S / 2 * H * D / t / 1e12:7.2f} Tflop/s causal", flush=True)
for S, H, D in [(577, 16, 64), (5476, 16, 80)]:
    q, k, v = (torch.randn((1, S, H, D), device="cuda") for _ in range(3))
    for mode in ("1", "0"):
        if mode == "0" and S > 2000:
            continue
        os.environ["MLLM_CUDA_ATTN_TC"] = mode
        t = timed(lambda i: capi.flash_attn_f32kv(q, k, v, causal=False), reps=3)
        print(f"attention fp32kv S={S} H={H} D={D} tc={mode}: {t * 1e6:8.1f} us  {4.0 * S * S * H * D / t / 1e12:7.2f} Tflop/s", flush=True)
