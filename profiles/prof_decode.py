"""Short profiling driver: load the 7B Q4_0 model, one prefill of 512 tokens, then N decode steps.
Used under ncu (launch list / --set full); never a source of bench numbers."""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
from mllm_b200 import backend, mllm_file as mf  # noqa: E402

prefill = int(sys.argv[1]) if len(sys.argv) > 1 else 512
decode = int(sys.argv[2]) if len(sys.argv) > 2 else 3
cfg = mf.config(bench.MODEL, cache_limit=bench.LIMIT)
path = f"/tmp/mllm_bench_{bench.MODEL}_q4_0.mllm"
bench.make_model_file(cfg, path)
m = backend.LlamaModel(cfg).load(path)
ids = np.random.default_rng(0).integers(0, cfg["vocab"], (1, prefill)).astype(np.float32)
a, b, t = m.request_device(ids, decode)
print("prefill ms", a, "decode ms", b, "launches", backend.kernel_launches(), "tokens", t.tolist())
