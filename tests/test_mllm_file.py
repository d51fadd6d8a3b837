"""CPU tests of the host logic: .mllm container round trip (mirrors test/quantizer/QuantTest.cpp
ReadTest/WriteTest), the random-init writer's Q4_0 bytes against the oracle, and the CPU model
runner (port vs compiled reference)."""
import numpy as np
import pytest

from mllm_b200 import mllm_file as mf
from oracle import pyoracle as po


def test_numpy_q4_0_writer_matches_oracle(oracle):
    rng = np.random.default_rng(3)
    w = (rng.standard_normal((37, 512)) * 0.02).astype(np.float32)
    w[0, :32] = 0
    w[1, 5] = -w[1, :32].max() * 2  # negative extreme -> positive d
    assert np.array_equal(mf.quantize_q4_0_np(w), oracle.quantize_q4_0(w))


def test_mllm_round_trip(tmp_path):
    cfg = mf.config("toy-gqa", cache_limit=64)
    t = mf.random_tensors(cfg, seed=5)
    path = str(tmp_path / "toy.mllm")
    mf.write_model(path, cfg, t)
    back, dts = mf.read_mllm(path)
    assert list(back) == [n for n, *_ in mf.tensor_list(cfg)]
    for name, rows, cols, dt in mf.tensor_list(cfg):
        assert dts[name] == dt
        assert back[name].nbytes == mf.nbytes(rows, cols, dt)
        assert np.array_equal(np.asarray(back[name]).view(np.uint8).ravel(), t[name].view(np.uint8).ravel())
    import os
    assert os.path.getsize(path) > mf.model_bytes(cfg)


def test_bad_magic_rejected(tmp_path):
    p = tmp_path / "bad.mllm"
    p.write_bytes(b"\x00" * 64)
    with pytest.raises(ValueError):
        mf.read_mllm(str(p))


def test_algorithmic_bytes_7b():
    cfg = mf.config("llama2-7b")
    # 6.607 G weights -> 3.717 GB per decoded token
    assert abs(mf.decode_weight_bytes(cfg) / 1e9 - 3.717) < 0.002


def test_cpu_model_runner_port_vs_reference(oracle, ref_lib):
    if ref_lib is None:
        pytest.skip("oracle/_ref not built here")
    cfg = mf.config("toy-gqa", cache_limit=32)
    t = mf.random_tensors(cfg, seed=9)
    names = mf.Names(cfg["rope_type"])
    rng = np.random.default_rng(1)
    ids = rng.integers(0, cfg["vocab"], (1, 8)).astype(np.float32)
    a = po.ModelRunner(oracle, cfg, t, names)
    b = po.ModelRunner(ref_lib, cfg, t, names)
    la, lb = a.forward(ids), b.forward(ids)
    assert np.abs(la - lb).max() <= 1e-4 * np.abs(lb).max()
    for _ in range(4):
        nxt = np.array([[la.argmax()]], dtype=np.float32)
        assert la.argmax() == lb.argmax()
        la, lb = a.forward(nxt), b.forward(nxt)
        assert np.abs(la - lb).max() <= 1e-4 * np.abs(lb).max()
