"""Timeline of the fused Linears inside real decode steps of the 7B model (development aid, MLLMCU_GEMV_PROF=1): the
%globaltimer stamps of CTA 0 for the last token's 129 launches -- entry, dependency wait passed, prologue done, exit."""
import ctypes as C
import os
import sys

import numpy as np

os.environ["MLLMCU_GEMV_PROF"] = "1"
os.environ["MLLMCU_ATTN_PROF"] = "1"
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
from mllm_b200 import backend, capi, mllm_file as mf  # noqa: E402

cfg = mf.config(bench.MODEL, cache_limit=bench.LIMIT)
path = f"/tmp/mllm_bench_{bench.MODEL}_q4_0.mllm"
bench.make_model_file(cfg, path)
m = backend.LlamaModel(cfg).load(path)
ids = np.random.default_rng(0).integers(0, cfg["vocab"], (1, 512)).astype(np.float32)
a, b, t = m.request_device(ids, 32)
print("prefill ms", a, "decode ms/token", b / 32)
n = 129
buf = np.zeros((n, 64), dtype=np.uint64)
got = capi.lib().mllmcu_gemv_prof_read(buf.ctypes.data_as(C.c_void_p), n)
assert got == n
ab = np.zeros((32, 8), dtype=np.uint64)
assert capi.lib().mllmcu_decode_attn_prof_read(ab.ctypes.data_as(C.c_void_p), 32) == 32
t0 = int(buf[0, 0])
names = ["qkv", "o", "gate|up", "down"]
tot = dict(gap=0.0, pro=0.0, stream=0.0)
for i in range(n):
    e = [(int(v) - t0) / 1e3 for v in buf[i, :4]]
    prev = (int(buf[i - 1, 3]) - t0) / 1e3 if i else 0.0
    nm = names[i % 4] if i < 128 else "lm_head"
    if 8 <= i < 16 or i >= 124:
        print(f"{nm:8s} entry {e[0]:8.2f} wait-done {e[1]:8.2f} prologue-done {e[2]:8.2f} exit {e[3]:8.2f} | prev exit -> wait-done {e[1] - prev:6.2f}  prologue {e[2] - e[1]:5.2f}  stream {e[3] - e[2]:5.2f}")
    if (8 <= i < 16 or i >= 124) and i % 4 == 0 and i < 128:
        a = [(int(v) - t0) / 1e3 for v in ab[i // 4]]
        print(f"  attn   entry {a[0]:8.2f} wait-done {a[1]:8.2f} q-rotated {a[2]:8.2f} keys-done {a[3]:8.2f} partial-stored {a[4]:8.2f} ticket {a[5]:8.2f}"
              f" | merge-start {a[6]:8.2f} out-stored {a[7]:8.2f}   (qkv exit {e[3]:8.2f})")
    if 4 <= i < 128:
        tot["gap"] += e[1] - prev
        tot["pro"] += e[2] - e[1]
        tot["stream"] += e[3] - e[2]
print("per layer (layers 1..31): previous exit -> wait-done %.2f us (attention sits in the qkv -> o gap), prologue %.2f us, stream %.2f us; total %.2f us"
      % (tot["gap"] / 31, tot["pro"] / 31, tot["stream"] / 31, sum(tot.values()) / 31))
