"""Parity on the BASELINE configs against the REFERENCE's own model graphs, run live on this box's host cores.

The reference side is oracle/_ref/libmllm_cpu_ref.so: the reference's CPU backend compiled from its own sources (oracle/cpuref.mk),
running TinyLLaMAModel / QWenForCausalLM / LLaMAModel on MLLM_CPU over the SAME random-init .mllm file the CUDA backend loads.

  C1  TinyLLaMA-1.1B Q4_0, prefill 128 + decode 128                     (BASELINE configs[0])
  C5  Qwen1.5-1.8B  Q4_0, batch 3 x 32-token prompts + 200 decode steps (configs[4], examples/demo_qwen_batch.cpp:61-68 shape)
  C2  LLaMA-2-7B    Q4_0, prefill 512 + decode 64                       (configs[1])

What can be asserted, and why (measured, profiles/r02_parity_baseline_configs.jsonl):
A deep Q8_0-activation model re-quantises its activations four times per block.  A rounding-level difference (1e-7 relative) in
any op flips a handful of those int8 roundings, each flip is a 1/127 step on one activation, and within a block the cascade saturates
at ~5e-3 of the block's output; through 22-32 blocks of a RANDOM-INIT model the logits of any two non-bit-identical
implementations end ~3-6e-2 apart.  The yardstick is the `floor` run: the reference's OWN compiled kernels with only the attention's
summation order changed (<= 1.2e-6 per call).  Its logits cosine against the reference is 0.99956 / 0.99963 / 0.99822 and its greedy
tokens leave the reference's after 6 / 6 / 1 steps -- so "identical tokens over 128 steps" is not a property even the reference has
against itself on these weights.  Asserted here:
  * the CPU runner over the reference's kernels IS the reference graph, bit for bit, at full size (logits and every token);
  * the CUDA logits are as close to the reference as the floor run is (1 - cos within 2.5x of the floor's);
  * CUDA greedy tokens equal the reference's until the first step whose reference top-2 margin is inside that noise (WHICH step that
    is moves with every change of summation order, for the floor run as for the CUDA path: 39 steps with one GEMV kernel, 2 with another);
  * every transformer block, fed the reference's own input of that block, reproduces the reference's output of that block
    (cosine >= 0.9999, 99.9 % of elements within 1e-2 of the block's max): no block is wrong, the differences are the flips.
Exact token identity over the full horizon IS asserted where it is a meaningful property: tests/test_model_gpu.py,
tests/test_decode_gpu.py (2-3 block models, 48-64 steps)."""
import json
import os

import numpy as np
import pytest

import parity_lib

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def be():
    import torch
    assert torch.cuda.is_available()
    from mllm_b200 import backend, build
    build.build_all()
    backend.init_backend(0)
    return backend


@pytest.fixture(scope="module")
def po():
    from oracle import pyoracle
    assert pyoracle.cpuref() is not None and pyoracle.ref() is not None, \
        "oracle/_ref/*.so missing: build them where /root/reference exists (make -C oracle); they travel with the tree"
    return pyoracle


def check(r):
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    with open(os.path.join(ROOT, "gpurun_out", "parity_baseline_configs.jsonl"), "a") as f:
        f.write(json.dumps({k: v for k, v in r.items() if k != "teacher_forced"}) + "\n")
    assert r["runner_equals_graph_bitwise"] and r["runner_tokens_equal_graph"]
    assert r["gpu_fused_decode"]
    floor, gpu = 1.0 - r["floor_cos_prefill"], 1.0 - r["gpu_cos_prefill"]
    assert gpu <= 2.5 * floor + 1e-6, (gpu, floor)
    assert r["gpu_relerr_prefill"] <= 2.5 * r["floor_relerr_prefill"] + 1e-6
    n = r["n_decode"]
    for b, k in enumerate(r["gpu_tokens_equal"]):
        if k == n:
            continue
        assert r["margin_at_first_mismatch"][b] <= r["gpu_relerr_prefill"], (b, k, r["margin_at_first_mismatch"][b])
    for t in r["teacher_forced"]:
        assert t["cos"] >= 0.9999, t
        assert t["p999"] <= 1e-2 and t["max"] <= 3e-2, t


def test_c1_tinyllama_128_plus_128(be, po):
    check(parity_lib.measure(be, po, "tinyllama-1.1b", 1, 128, 128, 512, seed=101))


def test_c5_qwen_batch3_32_plus_200(be, po):
    check(parity_lib.measure(be, po, "qwen1.5-1.8b", 3, 32, 200, 400, seed=105))


def test_c2_llama2_7b_512_plus_64(be, po):
    check(parity_lib.measure(be, po, "llama2-7b", 1, 512, 64, 1024, seed=102))
