"""Pins the CPU restatement (oracle/oracle.c, oracle/model_runner.c) against the REFERENCE's own Op classes and model graphs.

tests/golden/ref_ops.npz / ref_graphs.npz were produced by tests/golden/make_golden3.py from the reference's CPU backend compiled
in place (oracle/cpuref.mk): CPURoPE, CPURMSNorm, CPULayerNorm, CPUEmbedding, CPUMultimodalRoPE, CPUVisionRoPEFunc,
CPUConvolution2D/3D, CPUSoftMax, CPULinear created through CPUBackend::opCreate and driven reshape -> setUp -> execute as
test/cpu/CPUTest.hpp does; LLaMAModel / QWenForCausalLM run on MLLM_CPU over seeded .mllm files.  Every comparison is BIT-EXACT.
Where oracle/_ref/libmllm_cpu_ref.so is present (build container, GPU box) the same cases also run live."""
import os
import sys
import tempfile

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))
import refops_cases as rc  # noqa: E402
from make_golden3 import GRAPHS  # noqa: E402

from mllm_b200 import mllm_file as mf  # noqa: E402
from oracle import pyoracle as po  # noqa: E402


def bits(a):
    return np.ascontiguousarray(a, dtype=np.float32).view(np.uint32)


@pytest.fixture(scope="module")
def gold():
    return np.load(os.path.join(ROOT, "tests", "golden", "ref_ops.npz"))


@pytest.fixture(scope="module")
def gold_graphs():
    return np.load(os.path.join(ROOT, "tests", "golden", "ref_graphs.npz"))


@pytest.fixture(scope="module")
def all_cases(oracle):
    return rc.cases(oracle)


CASE_NAMES = ["rmsnorm", "rmsnorm_unit_offset", "rmsnorm_headwise", "layernorm", "layernorm_nobias", "rope_llama", "rope_hf", "rope_hf_theta1e6_d128",
              "rope_hf_partial", "rope_hf_llama3", "mrope", "vision_rope", "embedding_q4_0", "embedding_f32", "conv2d_patch", "conv3d_patch", "softmax",
              "linear_q4_0_bias", "linear_f32"]


@pytest.mark.parametrize("name", CASE_NAMES)
def test_restatement_equals_reference_op_bit_for_bit(oracle, gold, all_cases, name):
    case = all_cases[name]
    got = rc.run_port(oracle, name, case)
    if got is None:
        pytest.skip("no restatement for this op (golden used by the GPU tests directly)")
    assert len(got) == case.get("calls", 1)
    for k, y in enumerate(got):
        want = gold[f"{name}/{k}"]
        assert y.size == want.size
        y = y.reshape(want.shape)
        if "valid" in case:
            y, want = np.where(case["valid"], y, 0), np.where(case["valid"], want, 0)
        assert np.array_equal(bits(y), bits(want)), f"{name} call {k}: max abs diff {np.abs(y - want).max()}"


def test_case_list_is_complete(all_cases, gold):
    assert sorted(all_cases) == sorted(CASE_NAMES)
    assert sorted({k.split("/")[0] for k in gold.files}) == sorted(CASE_NAMES)


@pytest.mark.parametrize("name", CASE_NAMES)
def test_live_reference_op_equals_golden(gold, all_cases, name):
    if po.cpuref() is None:
        pytest.skip("oracle/_ref/libmllm_cpu_ref.so not built here")
    for k, y in enumerate(rc.run_reference(po, all_cases[name])):
        want = gold[f"{name}/{k}"]
        if "valid" in all_cases[name]:
            y, want = np.where(all_cases[name]["valid"], y, 0), np.where(all_cases[name]["valid"], want, 0)
        assert np.array_equal(bits(y), bits(want))


@pytest.mark.parametrize("name,B,S,steps", GRAPHS)
def test_model_runner_equals_reference_graph(oracle, gold_graphs, name, B, S, steps):
    """oracle/model_runner.c over the restatement reproduces the logits of the reference's own model graph bit for bit, prefill and
    every greedy decode step (so RoPE positions, KV-cache append, fp16 K/V rounding, GQA head mapping, biases and the clip to the last
    position all agree with CPURoPE / CPUKVCache / MultiHeadAttention / LLaMAModel / QWenForCausalLM)."""
    cfg = mf.config(name, cache_limit=64)
    t = mf.random_tensors(cfg, seed=1)
    run = po.ModelRunner(oracle, cfg, t, mf.Names(cfg["rope_type"]), max_batch=B)
    ids = gold_graphs[f"{name}/ids"]
    want_logits, want_tok = gold_graphs[f"{name}/logits"], gold_graphs[f"{name}/tokens"]
    lg = run.forward(ids)
    assert np.array_equal(bits(lg), bits(want_logits[0]))
    for k in range(steps):
        nxt = lg.argmax(axis=1)
        assert np.array_equal(nxt.astype(np.int32), want_tok[k])
        lg = run.forward(nxt.astype(np.float32).reshape(B, 1))
        assert np.array_equal(bits(lg), bits(want_logits[k + 1])), f"decode step {k}"


def test_live_reference_graph_equals_golden(gold_graphs):
    if po.cpuref() is None:
        pytest.skip("oracle/_ref/libmllm_cpu_ref.so not built here")
    name, B, S, steps = GRAPHS[1]
    cfg = mf.config(name, cache_limit=64)
    t = mf.random_tensors(cfg, seed=1)
    with tempfile.TemporaryDirectory() as d:
        path = os.path.join(d, "m.mllm")
        mf.write_model(path, cfg, t)
        g = po.RefGraph(cfg, path, threads=2)
        lg = g.forward(gold_graphs[f"{name}/ids"])
        assert np.array_equal(bits(lg), bits(gold_graphs[f"{name}/logits"][0]))
        out, n = g.generate(gold_graphs[f"{name}/ids"], 4)   # a second request without clear: Module::generate continues the cache
        g.close()
