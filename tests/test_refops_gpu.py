"""CUDA ops through Backend::opCreate + runOp (the test/cpu/CPUTest.hpp per-op pattern) against outputs of the REFERENCE's own CPU Op
classes: tests/golden/ref_ops.npz, written by tests/golden/make_golden3.py from oracle/_ref/libmllm_cpu_ref.so.  No restatement in
between: the tolerance of each op is stated next to it."""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))
import refops_cases as rc  # noqa: E402

from mllm_b200 import mllm_file as mf  # noqa: E402

pytestmark = pytest.mark.gpu

# case -> (relative tolerance against max|want|; 0 = bit-exact)
TOL = {
    "rmsnorm": 2e-7, "rmsnorm_unit_offset": 2e-7, "rmsnorm_headwise": 2e-7,   # <= 1 ulp: fp64 sum of squares in tree order, same float mean
    "layernorm": 2e-6, "layernorm_nobias": 2e-6,                              # the two fp32 sums are tree-ordered here, serial there
    "rope_llama": 0, "rope_hf": 0, "rope_hf_theta1e6_d128": 0, "rope_hf_partial": 0, "rope_hf_llama3": 0,  # host-built tables, FMA form
    "mrope": 0,
    "vision_rope": 3e-6,   # the op evaluates sinf / cosf on the device (<= 2 ulp from libm); with host-built tables the kernel is bit-exact (tests/test_vision_kq_gpu.py)
    "embedding_q4_0": 0, "embedding_f32": 0,
    "linear_f32": 1e-5,                                                       # F32-weight Linear: vec_dot_fp32's 32-lane order there, tiled fp32 sums here
    "conv2d_patch": 1e-5,                                                     # vec_dot_fp32's 32-lane order there, tiled sums here
    "softmax": 0,
    "linear_q4_0_bias": 2e-5,                                                 # same products, different fp32 order (1e-3)
}


@pytest.fixture(scope="module")
def be():
    import torch
    assert torch.cuda.is_available()
    from mllm_b200 import backend, build
    build.build_all()
    backend.init_backend(0)
    return backend


@pytest.fixture(scope="module")
def gold():
    return np.load(os.path.join(ROOT, "tests", "golden", "ref_ops.npz"))


@pytest.fixture(scope="module")
def all_cases(oracle):
    return rc.cases(oracle)


@pytest.mark.parametrize("name", sorted(TOL))
def test_cuda_op_against_reference_op(be, gold, all_cases, tmp_path, name):
    case = all_cases[name]
    path = ""
    if case["weights"]:
        path = str(tmp_path / "w.mllm")
        mf.write_mllm(path, {"op" + s: np.ascontiguousarray(a) for s, (a, _) in case["weights"].items()}, {"op" + s: dt for s, (_, dt) in case["weights"].items()})
    ins = [np.ascontiguousarray(a, dtype=np.float32) for a, *_ in case["inputs"]]
    shapes = [list(shape) for _, shape, *_ in case["inputs"]]
    y = be.run_op(case["params"], "op", ins, path=path, shapes=shapes)
    want = gold[f"{name}/0"]
    y = y.reshape(want.shape)
    if "valid" in case:
        y, want = np.where(case["valid"], y, 0), np.where(case["valid"], want, 0)
    if TOL[name] == 0:
        assert np.array_equal(y.view(np.uint32), want.view(np.uint32)), f"max abs diff {np.abs(y - want).max()}"
    else:
        assert np.abs(y - want).max() <= TOL[name] * np.abs(want).max(), (np.abs(y - want).max() / np.abs(want).max(), int((y.view(np.uint32) != want.view(np.uint32)).sum()), y.size)
