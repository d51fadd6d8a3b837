"""Regenerates tests/golden/ref_ops.npz and ref_graphs.npz from the REFERENCE's own CPU backend compiled in place
(oracle/cpuref.mk -> oracle/_ref/libmllm_cpu_ref.so): per-op outputs of its Op classes created the way test/cpu/CPUTest.hpp does,
and logits / greedy tokens of its model graphs (LLaMAModel, QWenForCausalLM) on MLLM_CPU over seeded random-init .mllm files.

    python tests/golden/make_golden3.py        (needs /root/reference; run in the build container)
"""
import os
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))

from mllm_b200 import mllm_file as mf  # noqa: E402  (file-format writer only)
from oracle import pyoracle as po  # noqa: E402
import refops_cases as rc  # noqa: E402

# prompt lengths are multiples of 4: for Sk % 4 != 0 the reference's fp16-KV prefill mis-sizes its last key tile
# (compute/FlashAttention2.hpp:1274 uses `% Tc` where `% Bc` is meant), decode is unaffected
GRAPHS = [("toy-llama", 1, 8, 12), ("toy-gqa", 1, 8, 12), ("toy-qwen", 3, 8, 10), ("toy-d128", 2, 4, 8)]


def main():
    po.build()
    po.build_cpuref()
    P = po.port()
    out = {}
    for name, case in rc.cases(P).items():
        for k, y in enumerate(rc.run_reference(po, case)):
            out[f"{name}/{k}"] = y
    np.savez_compressed(os.path.join(ROOT, "tests", "golden", "ref_ops.npz"), **out)
    g = {}
    for name, B, S, steps in GRAPHS:
        cfg = mf.config(name, cache_limit=64)
        t = mf.random_tensors(cfg, seed=1)
        with tempfile.TemporaryDirectory() as d:
            path = os.path.join(d, "m.mllm")
            mf.write_model(path, cfg, t)
            graph = po.RefGraph(cfg, path, threads=4)
            ids = np.random.default_rng(11).integers(0, cfg["vocab"], (B, S)).astype(np.float32)
            logits = [graph.forward(ids)]
            toks = []
            for _ in range(steps):
                nxt = logits[-1].argmax(axis=1).astype(np.float32)
                toks.append(nxt.astype(np.int32))
                logits.append(graph.forward(nxt.reshape(B, 1)))
            g[f"{name}/ids"] = ids
            g[f"{name}/logits"] = np.stack(logits)
            g[f"{name}/tokens"] = np.stack(toks)
            graph.close()
    np.savez_compressed(os.path.join(ROOT, "tests", "golden", "ref_graphs.npz"), **g)
    print("wrote", len(out), "op outputs and", len(GRAPHS), "graphs")


if __name__ == "__main__":
    main()
