"""Compare two MLLM_DUMP_OPS directories (integration built with DEBUG_DUMP=1) op by op: argv = cpu_dir cuda_dir."""
import os
import sys

import numpy as np


def load(p):
    raw = np.fromfile(p, dtype=np.uint8)
    b, h, s, d = raw[:16].view(np.int32)
    return raw[16:].view(np.float32).reshape(b, s, h, d)


a_dir, b_dir = sys.argv[1:3]
names = sorted(os.listdir(a_dir), key=lambda n: os.path.getmtime(os.path.join(a_dir, n)))
for n in names:
    pb = os.path.join(b_dir, n)
    if not os.path.exists(pb):
        print(f"{n:70s} only in {a_dir}")
        continue
    a, b = load(os.path.join(a_dir, n)), load(pb)
    if a.shape != b.shape:
        print(f"{n:70s} shapes {a.shape} vs {b.shape}")
        continue
    m = np.abs(a).max() or 1.0
    err = np.abs(a - b).max() / m
    print(f"{n:70s} {str(a.shape):22s} rel max err {err:9.2e}  max|a| {m:9.3g}")
