"""CPU tests of the drop-in boundary: libmllmcu.so builds for sm_100a, loads without a
GPU, and exports every symbol include/mllmcu.h declares (no compute calls here)."""
import ctypes
import os
import re

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def header_symbols():
    src = open(os.path.join(ROOT, "include", "mllmcu.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(mllmcu_[A-Za-z0-9_]+)\s*\(", src)))


def test_library_builds_and_exports_header():
    from mllm_b200 import build, capi
    so = build.build_kernels()
    assert os.path.exists(so)
    L = ctypes.CDLL(so)
    syms = header_symbols()
    assert len(syms) >= 25
    for s in syms:
        assert hasattr(L, s), f"{s} declared in include/mllmcu.h but not exported"
    # the ctypes table binds exactly the header's symbols
    assert sorted(capi.SYMBOLS) == syms


def test_version_and_no_gpu_is_loud():
    from mllm_b200 import capi
    L = capi.lib()
    assert L.mllmcu_version() >= 100
    import torch
    if not torch.cuda.is_available():
        # no device: memory calls must fail with an error code, never silently succeed
        p = ctypes.c_void_p()
        assert L.mllmcu_malloc(ctypes.byref(p), 1024) != 0


def test_product_never_imports_oracle():
    """The product package must not reference the CPU checkers."""
    bad = []
    for dp, _, fns in os.walk(os.path.join(ROOT, "mllm_b200")):
        for fn in fns:
            if fn.endswith((".py", ".cu", ".cuh", ".cpp", ".hpp", ".h")):
                txt = open(os.path.join(dp, fn), errors="ignore").read()
                if re.search(r"pyoracle|liboracle|libmllm_ref|oracle/|orc_|from oracle|import oracle", txt):
                    bad.append(os.path.join(dp, fn))
    assert not bad, bad
