"""fp32-K/V (vision tower) attention timed alone: Qwen2-VL-2B tower (5476 patches, 16 heads x 80) and CLIP ViT-L/14-336 (577 x 16 x 64);
K / V pre-split into fp16 planes (default workspace) against the in-kernel split (small workspace).  CUDA events, 10 launches after 3 warm-ups;
and the patch-embed convolution of the Qwen2-VL processor layout (5476 entries of 3 x 2 x 14 x 14 -> 1280).
usage: PYTHONPATH=. python profiles/bench_vit_attn.py"""
import json
import torch
from mllm_b200 import capi as cu

dev = torch.device("cuda:0")


def timed(fn, n=10):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


for name, S, H, D in (("qwen2vl tower", 5476, 16, 80), ("clip vit-l", 577, 16, 64)):
    q, k, v = (torch.randn((1, S, H, D), device=dev) for _ in range(3))
    flop = 4.0 * S * S * H * D
    out = {"attention": name, "S": S, "H": H, "D": D}
    for label, small in (("planes", False), ("in_kernel_split", True)):
        ms = timed(lambda: cu.flash_attn_f32kv(q, k, v, causal=False, small_ws=small))
        out[label + "_ms"] = round(ms, 4)
        out[label + "_TFLOPs"] = round(flop / ms / 1e9, 1)
    print(json.dumps(out), flush=True)

x = torch.randn((5476, 3, 2, 14, 14), device=dev)
w = torch.randn((1280, 3, 2, 14, 14), device=dev) * 0.02
b = torch.randn((1280,), device=dev)
ms = timed(lambda: cu.conv3d_valid(x, w, b, (2, 14, 14), channels_last=True))
print(json.dumps({"conv3d": "qwen2vl patch embed 5476 x 1176 -> 1280", "ms": round(ms, 4), "TFLOPs": round(2.0 * 5476 * 1176 * 1280 / ms / 1e9, 2)}))
