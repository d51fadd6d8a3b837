"""Q4_K decode GEMV kernels timed alone (CUDA events around a captured graph of launches over rotating weight copies larger than L2):
mllmcu_gemv_q4k on native and on 304-byte activation blocks, and the op-level Linear (quantise + k_gemv_kq), on the LLaVA-1.5-7B shapes.
(r02_q4k_gemv_micro_v1.jsonl: the first version -- per-super-block shuffles, "direct" -- and a TMA-ring variant, "ring", which was slower.)
usage: python profiles/bench_q4k_gemv.py"""
import ctypes as C
import json
import numpy as np
import torch
from mllm_b200 import capi as cu

dev = torch.device("cuda:0")
rng = np.random.default_rng(0)
SHAPES = [("o 4096x4096", (4096,), 4096), ("qkv 3x4096x4096", (4096, 4096, 4096), 4096), ("gate|up 2x11008x4096", (11008, 11008), 4096),
          ("down 4096x11008", (4096,), 11008), ("lm_head 32000x4096", (32000,), 4096)]


def rand_q4k(N, K):
    W = torch.randint(0, 256, (N, K // 256, 144), dtype=torch.uint8, device=dev)
    d = torch.full((N, K // 256, 1), 0.01, dtype=torch.float16, device=dev).view(torch.uint8)
    W[:, :, 0:2] = d
    W[:, :, 2:4] = d
    return W.reshape(N, -1)


def timed(fn, reps):
    g = torch.cuda.CUDAGraph()
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        fn(0)
        torch.cuda.synchronize()
        with torch.cuda.graph(g, stream=s):
            for r in range(reps):
                fn(r)
    for _ in range(3):
        g.replay()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5):
        g.replay()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) * 1000 / (5 * reps)


for M in (1, 8):
    for name, Ns, K in SHAPES:
        bytes_w = sum(Ns) * (K // 256) * 144
        copies = max(2, int(300e6 // bytes_w) + 1)
        sets = [[rand_q4k(N, K) for N in Ns] for _ in range(copies)]
        x = torch.randn((M, K), device=dev)
        xqs = {1: cu.prep_q8_K(x), 3: cu.prep_q8_K(x, aligned=True)}
        xqs[7] = xqs[3]
        ys = [torch.empty((M, N), device=dev) for N in Ns]
        n = len(Ns)
        Np = (C.c_int * n)(*Ns)
        yp = (C.c_void_p * n)(*[y.data_ptr() for y in ys])
        Wps = [(C.c_void_p * n)(*[w.data_ptr() for w in ws]) for ws in sets]
        bp = (C.c_void_p * n)(*[None] * n)
        reps = copies * 2
        out = {"M": M, "shape": name, "MB": round(bytes_w / 1e6, 1)}
        for label, flags in (("native", 1), ("aligned", 3), ("aligned_no_areg", 7)):
            def fn(r, flags=flags):
                st = torch.cuda.current_stream().cuda_stream
                # all segments share ldy only when equal N; launch per distinct ldy otherwise -- here every segment has the same N
                cu.check(cu.lib().mllmcu_gemv_q4k(n, Wps[r % copies], Np, bp, yp, Ns[0], xqs[flags].data_ptr(), M, K, None, 0, flags, st), "gemv_q4k")
            us = timed(fn, reps)
            out[label + "_us"] = round(us, 2)
            out[label + "_TBps"] = round(bytes_w / us / 1e6, 2)
        def fn_old(r):
            for W, y in zip(sets[r % copies], ys):
                cu.linear(W, x, wtype=cu.TYPE_Q4_K, out=y, ldy=y.shape[1])
        us = timed(fn_old, reps)
        out["linear_us"] = round(us, 2)
        out["linear_TBps"] = round(bytes_w / us / 1e6, 2)
        print(json.dumps(out), flush=True)
        del sets
        torch.cuda.empty_cache()
