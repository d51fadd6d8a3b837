"""Model-level parity measurements shared by tests/test_baseline_configs_gpu.py and profiles/parity_baseline_configs.py.

Three CPU runs over the same .mllm file, all on the reference's arithmetic:
  graph   the reference's own model graph on MLLM_CPU (oracle/_ref/libmllm_cpu_ref.so) -- the truth for logits and tokens;
  runner  oracle/model_runner.c over the reference's compiled kernels (bit-identical to `graph`, tested) -- gives every layer's input;
  floor   the same runner with ONLY the attention's summation order changed (restatement order, <= 1.2e-6 per call).
`floor` measures how far a rounding-level difference in one op travels through the Q8_0 re-quantisations of a deep random-init model:
it is what ANY backend that is not bit-identical in every op looks like from the logits, and the yardstick for the CUDA path."""
import os

import numpy as np

from mllm_b200 import mllm_file as mf


def host_threads():
    n = len(os.sched_getaffinity(0))
    return max(1, min(32, n // 2))   # SMT siblings / oversubscribed vCPUs stall the reference's spin-waiting OpenMP teams


def cosine(a, b):
    a, b = a.astype(np.float64).ravel(), b.astype(np.float64).ravel()
    return float((a * b).sum() / (np.linalg.norm(a) * np.linalg.norm(b) + 1e-30))


def greedy(model, ids, n):
    """n greedy tokens from any object with forward(ids)->[B,V] (prefill + n-1 single-token steps); also the top-2 margins."""
    toks, margins, cur = [], [], ids
    for _ in range(n):
        lg = model.forward(cur)
        top2 = np.sort(lg, axis=1)[:, -2:]
        margins.append((top2[:, 1] - top2[:, 0]) / np.abs(lg).max(axis=1))
        toks.append(lg.argmax(axis=1))
        cur = toks[-1].astype(np.float32)[:, None]
    return np.stack(toks, axis=1).astype(np.int32), np.stack(margins, axis=1)


def measure(be, po, name, B, S, n_decode, limit, seed, layers_checked=None):
    cfg = mf.config(name, cache_limit=limit)
    path = f"/tmp/mllm_test_{name}_q4_0.mllm"
    mf.write_random_model_torch(cfg, path, device="cuda", seed=1234)
    ids = np.random.default_rng(seed).integers(0, cfg["vocab"], (B, S)).astype(np.float32)
    th = host_threads()
    os.environ["OMP_NUM_THREADS"] = str(th)
    out = {"config": name, "B": B, "S": S, "n_decode": n_decode, "host_threads": th}
    # ---- truth: the reference graph
    graph = po.RefGraph(cfg, path, threads=th)
    lr0 = graph.forward(ids)
    graph.reset()
    want, n = graph.generate(ids, n_decode)   # Module::generate, greedy (Module.cpp:146-205)
    want = want.astype(np.int32)
    graph.close()
    assert (n == n_decode).all()
    # ---- the runner over the reference's kernels: layer inputs, margins, and the rounding-noise floor
    tensors, _ = mf.read_mllm(path)
    runner = po.ModelRunner(po.ref(), cfg, tensors, mf.Names(cfg["rope_type"]), max_batch=B)
    lc0, dump = runner.forward(ids, dump=True)
    out["runner_equals_graph_bitwise"] = bool(np.array_equal(lc0.view(np.uint32), lr0.view(np.uint32)))
    runner.reset()
    rt, margins = greedy(runner, ids, n_decode)
    out["runner_tokens_equal_graph"] = bool(np.array_equal(rt, want))
    runner.reset()
    runner.set_port_attention(True)
    lf0 = runner.forward(ids)
    runner.reset()
    ft, _ = greedy(runner, ids, n_decode)
    runner.set_port_attention(False)
    out["floor_cos_prefill"] = cosine(lf0, lr0)
    out["floor_tokens_equal"] = [int(np.argmin(r)) if not r.all() else n_decode for r in (ft == want)]
    # ---- CUDA backend
    gpu = be.LlamaModel(cfg).load(path)
    lg0 = gpu.forward(ids)
    gpu.clear_kvcache()
    toks = gpu.generate(ids, n_decode)
    out["gpu_fused_decode"] = bool(gpu.last_step_fused())
    out["gpu_cos_prefill"] = cosine(lg0, lr0)
    out["gpu_relerr_prefill"] = float(np.abs(lg0 - lr0).max() / np.abs(lr0).max())
    out["floor_relerr_prefill"] = float(np.abs(lf0 - lr0).max() / np.abs(lr0).max())
    same = toks == want
    out["gpu_tokens_equal"] = [int(np.argmin(r)) if not r.all() else n_decode for r in same]
    out["margin_at_first_mismatch"] = [float(margins[b, k]) if k < n_decode else None for b, k in enumerate(out["gpu_tokens_equal"])]
    out["min_margin_before_mismatch"] = [float(margins[b, :max(k, 1)].min()) for b, k in enumerate(out["gpu_tokens_equal"])]
    # ---- teacher-forced blocks: the reference's own layer input in, compare that layer's output
    L = cfg["layers"]
    H = cfg["hidden"]
    tf = []
    for layer in (range(L) if layers_checked is None else layers_checked):
        x = dump[layer].reshape(B, S, H)
        y = gpu.block_forward(layer, x).reshape(-1, H)
        ref = dump[layer + 1]
        err = np.abs(y - ref)
        scale = float(np.abs(ref).max())
        tf.append({"layer": int(layer), "max": float(err.max() / scale), "p999": float(np.quantile(err, 0.999) / scale),
                   "over_1e-3": int((err > 1e-3 * scale).sum()), "n": int(err.size), "cos": cosine(y, ref)})
    out["teacher_forced"] = tf
    gpu.close()
    return out
