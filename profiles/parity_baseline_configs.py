"""python profiles/parity_baseline_configs.py [c1] [c5] [c2]  -- one JSON line per config (see tests/parity_lib.py)."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import parity_lib  # noqa: E402

from mllm_b200 import backend, build  # noqa: E402
from oracle import pyoracle as po  # noqa: E402

CASES = {"c1": ("tinyllama-1.1b", 1, 128, 128, 512, 101), "c5": ("qwen1.5-1.8b", 3, 32, 200, 400, 105), "c2": ("llama2-7b", 1, 512, 64, 1024, 102)}
build.build_all()
backend.init_backend(0)
for k in (sys.argv[1:] or list(CASES)):
    r = parity_lib.measure(backend, po, *CASES[k])
    tf = r.pop("teacher_forced")
    r["teacher_forced_worst"] = max(tf, key=lambda d: d["max"])
    r["teacher_forced_over_1e-3_total"] = sum(d["over_1e-3"] for d in tf)
    r["teacher_forced_min_cos"] = min(d["cos"] for d in tf)
    print(json.dumps({"case": k, **r}), flush=True)
