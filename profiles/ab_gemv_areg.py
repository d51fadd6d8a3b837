import os, sys, numpy as np
sys.path.insert(0, os.getcwd())
import bench
from mllm_b200 import backend, mllm_file as mf
cfg = mf.config(bench.MODEL, cache_limit=bench.LIMIT)
path = f"/tmp/mllm_bench_{bench.MODEL}_q4_0.mllm"
bench.make_model_file(cfg, path)
m = backend.LlamaModel(cfg).load(path)
ids = np.random.default_rng(0).integers(0, cfg["vocab"], (1, 512)).astype(np.float32)
for r in range(3):
    m.clear_kvcache()
    a, b, t = m.request_device(ids, 256)
    print("areg", os.environ.get("MLLMCU_GEMV_AREG", "1"), "prefill ms %.2f decode ms/token %.4f" % (a, b / 255), "tokens", t.tolist()[0][:8], flush=True)
