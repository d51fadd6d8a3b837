#!/usr/bin/env python
"""Decode speed with Q4_K Linears (the LLaVA-1.5 LLM's dtype, BASELINE configs[2]) on a TinyLLaMA-sized model: the fused decode step
covers Q4_0 only, so this runs op by op.  python profiles/bench_kquant_decode.py [--model tinyllama-1.1b]"""
import argparse
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from mllm_b200 import backend, mllm_file as mf  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--model", default="tinyllama-1.1b")
ap.add_argument("--prompt", type=int, default=128)
ap.add_argument("--decode", type=int, default=64)
a = ap.parse_args()
cfg = mf.config(a.model, cache_limit=a.prompt + a.decode + 8)
cfg["wtype"] = mf.Q4_K
path = f"/tmp/{a.model}-q4_k.mllm"
if not os.path.exists(path):
    t0 = time.time()
    mf.write_model(path, cfg, mf.random_tensors(cfg, seed=3))
    print(f"model file written in {time.time() - t0:.0f}s", file=sys.stderr)
model = backend.LlamaModel(cfg).load(path)
ids = np.random.default_rng(1).integers(0, cfg["vocab"], (1, a.prompt)).astype(np.float32)
for _ in range(2):
    model.clear_kvcache()
    ms_p, ms_d, _ = model.request_device(ids, a.decode)
print(f"{a.model} Q4_K: prefill {a.prompt} tokens {ms_p:.1f} ms, decode {ms_d / a.decode:.3f} ms/token ({1e3 * a.decode / ms_d:.0f} tok/s), decode_mode {model.decode_mode()}")
