"""Optional tensor-parallel path on GPUs (mllm_b200/tp.py, csrc/p2p.cu).
1 GPU: the TP code with world = 1 (no exchange) must decode exactly what the product's own LlamaModel decodes -- same kernels, same order.
2 GPUs (skipped on a 1-GPU box; run with `gpurun --gpus 2`): logits of the sharded run against the 1-GPU run within the fp32
partial-sum tolerance, same tokens, for both exchange paths (peer-memory kernel, NCCL)."""
import json
import os
import subprocess
import sys

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
RUN = os.path.join(ROOT, "profiles", "tp_run.py")


def launch(nproc, args, env=None, port=29650):
    cmd = [sys.executable, RUN] if nproc == 1 else [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={nproc}", "--master-addr", "127.0.0.1",
                                                     "--master-port", str(port), RUN]
    r = subprocess.run(cmd + [str(x) for x in args], capture_output=True, text=True, timeout=900, env=dict(os.environ, **(env or {})))
    assert r.returncode == 0, r.stderr[-3000:]
    return json.loads([l for l in r.stdout.splitlines() if l.startswith("{")][-1])


@pytest.mark.parametrize("model", ["toy-gqa", "toy-llama"])
def test_world_1_equals_the_product_model(model, tmp_path):
    from mllm_b200 import backend, build, mllm_file as mf
    build.build_all()
    backend.init_backend(0)
    one = launch(1, ["--model", model, "--prompt", 16, "--decode", 24, "--limit", 128, "--save", tmp_path / "one.npz"])
    cfg = mf.config(model, cache_limit=128)
    path = str(tmp_path / "m.mllm")
    mf.write_model(path, cfg, mf.random_tensors(cfg, seed=1234))
    ids = np.random.default_rng(1234).integers(0, cfg["vocab"], size=(1, 16)).astype(np.float32)
    gpu = backend.LlamaModel(cfg).load(path)
    want = gpu.generate(ids, 25)[0].tolist()
    gpu.close()
    assert one["tokens"][:25] == want
    assert one["launches_per_step"] == 3 + 4 * cfg["layers"]


@pytest.mark.parametrize("exchange", ["p2p-fused-in-gemv", "p2p-kernel", "nccl"])
def test_two_ranks_match_one(exchange, tmp_path):
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs (gpurun --gpus 2)")
    args = ["--model", "toy-gqa", "--prompt", 16, "--decode", 24, "--limit", 128]
    one = launch(1, args + ["--save", tmp_path / "one.npz"])
    two = launch(2, args + ["--save", tmp_path / "two.npz"], env={"MLLM_TP_NCCL": "1" if exchange == "nccl" else "0", "MLLM_TP_FUSED": "1" if exchange == "p2p-fused-in-gemv" else "0"})
    assert two["tp"] == 2 and two["exchange"] == exchange
    a, b = np.load(tmp_path / "one.npz"), np.load(tmp_path / "two.npz")
    for key in ("logits0",):  # the prefill logits: before any token feedback
        err = np.abs(a[key] - b[key]).max() / np.abs(a[key]).max()
        assert err < 2e-2, (key, err)  # o / down partial sums meet in another order; later layers re-quantise (tests/test_model_gpu.py bars)
    k = next((i for i, (x, y) in enumerate(zip(one["tokens"], two["tokens"])) if x != y), len(one["tokens"]))
    assert k >= 4, (one["tokens"], two["tokens"])
