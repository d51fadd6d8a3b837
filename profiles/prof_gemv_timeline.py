"""Timeline of PDL-chained decode GEMV launches inside one CUDA graph (development aid, MLLMCU_GEMV_PROF=1):
per launch the %globaltimer stamps of CTA 0 -- entry, dependency wait passed, prologue done, exit."""
import ctypes as C
import os
import sys

import numpy as np
import torch

os.environ["MLLMCU_GEMV_PROF"] = "1"
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from mllm_b200 import capi  # noqa: E402

H, F = 4096, 11008
mk = lambda n, k: torch.randint(0, 255, (n, k // 32 * 18), dtype=torch.uint8, device="cuda")
NL = 12
layers = [dict(q=mk(H, H), k=mk(H, H), v=mk(H, H), o=mk(H, H), g=mk(F, H), u=mk(F, H), d=mk(H, F)) for _ in range(NL)]
x = torch.randn((1, H), device="cuda")
nw = torch.ones(H, device="cuda")


LEAD = int(sys.argv[1]) << 20 if len(sys.argv) > 1 else 0
own = []
for L in layers:
    own += [[L["q"], L["k"], L["v"]], [L["o"]], [L["g"], L["u"]], [L["d"]]]


def lookahead():
    """same plan as build_lookahead() in csrc/host/Module.cpp"""
    segs, ends, total = [], [], 0
    for launch in own:
        for w in launch:
            segs.append((w.data_ptr(), w.numel(), total))
            total += w.numel()
        ends.append(total)
    out, frontier, si = [], LEAD, 0
    for i in range(len(own)):
        la = []
        while LEAD and frontier < ends[i] + LEAD and len(la) < 8:
            pos = frontier % total
            while not (segs[si][2] <= pos < segs[si][2] + segs[si][1]):
                si = (si + 1) % len(segs)
            off = pos - segs[si][2]
            take = min(segs[si][1] - off, ends[i] + LEAD - frontier)
            la.append((segs[si][0] + off, take))
            frontier += take
        out.append(la)
    return out


PLAN = lookahead()


R = int(sys.argv[2]) if len(sys.argv) > 2 else 1
nwr = nw.repeat(R, 1).contiguous()
QKV = 3 * H
bufs = dict(qkv=torch.empty((1, 1, QKV), device="cuda"), x=torch.randn((R, 1, H), device="cuda"), x2=torch.empty((R, 1, H), device="cuda"),
            g=torch.empty((R, 1, F), device="cuda"))
bufs["x"][:] = bufs["x"][0]


SAME = len(sys.argv) > 3 and sys.argv[3] == "same"  # every launch the same kernel variant (o-type): is the prologue I-cache bound?


def step():
    if SAME:
        for li, L in enumerate(layers):
            for j, w in enumerate(("q", "k", "v", "o")):
                src, dst = (bufs["x"], bufs["x2"]) if j % 2 == 0 else (bufs["x2"], bufs["x"])
                capi.gemv_fused_ex([L[w]], src, dst, epilogue=1, residual=src[0], pdl=True, x_reps=R, y_reps=R)
        return
    for li, L in enumerate(layers):
        capi.gemv_fused_ex([L["q"], L["k"], L["v"]], bufs["x"], bufs["qkv"], norm_w=nwr, pdl=True, x_reps=R, norm_reps=R, lookahead=PLAN[4 * li])
        capi.gemv_fused_ex([L["o"]], bufs["qkv"][:, :, :H], bufs["x2"], epilogue=1, residual=bufs["x"][0], pdl=True, y_reps=R, lookahead=PLAN[4 * li + 1])
        capi.gemv_fused_ex([L["g"], L["u"]], bufs["x2"], bufs["g"], norm_w=nwr, epilogue=2, pdl=True, x_reps=R, norm_reps=R, y_reps=R,
                           lookahead=PLAN[4 * li + 2])
        capi.gemv_fused_ex([L["d"]], bufs["g"], bufs["x"], epilogue=1, residual=bufs["x2"][0], pdl=True, x_reps=R, y_reps=R, lookahead=PLAN[4 * li + 3])


for _ in range(2):
    step()
side = torch.cuda.Stream()
side.wait_stream(torch.cuda.current_stream())
with torch.cuda.stream(side):
    gr = torch.cuda.CUDAGraph()
    with torch.cuda.graph(gr, stream=side):
        step()
    gr.replay()
    gr.replay()
torch.cuda.synchronize()
n = 4 * len(layers)
buf = np.zeros((n, 64), dtype=np.uint64)
got = capi.lib().mllmcu_gemv_prof_read(buf.ctypes.data_as(C.c_void_p), n)
t0 = buf[0, 0]
names = ["qkv", "o", "gate|up", "down"]
ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
with torch.cuda.stream(side):
    ev[0].record(side)
    for _ in range(20):
        gr.replay()
    ev[1].record(side)
torch.cuda.synchronize()
print(f"lead {LEAD >> 20} MB, {R} replicas: {ev[0].elapsed_time(ev[1]) / 20 / NL * 1e3:.2f} us per layer (4 launches), graph of {NL} layers")
got = min(got, 16)
print("launch      entry   wait-done  prologue-done  exit   (us since first entry)   | gap from previous exit to entry, to wait-done")
for i in range(got):
    e = [(int(v) - int(t0)) / 1e3 for v in buf[i]]
    prev = (int(buf[i - 1, 3]) - int(t0)) / 1e3 if i else 0.0
    print(f"{names[i % 4]:8s} {e[0]:9.2f} {e[1]:9.2f} {e[2]:9.2f} {e[3]:9.2f}   | {e[0] - prev:6.2f} {e[1] - prev:6.2f}   prologue {e[2] - e[1]:5.2f}  stream {e[3] - e[2]:5.2f}")
    if 4 <= i < 8:
        print("      per warp loads landed/quantisation done, us after wait:", " ".join(f"{e[8 + 2 * w] - e[1]:.2f}/{e[9 + 2 * w] - e[1]:.2f}" for w in range(16) if buf[i, 9 + 2 * w]),
              f" rms known {e[4] - e[1]:.2f}" if buf[i, 4] else "")
