#!/usr/bin/env python
"""BASELINE configs[4]: Qwen1.5-1.8B Q4_0 batched greedy decode (examples/demo_qwen_batch.cpp shape: B prompts of 32 tokens,
200 decode steps), one replica per GPU, requests sharded data-parallel with no collective.

    python profiles/bench_batch_decode.py [--batch 8] [--prompt 32] [--decode 200]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P profiles/bench_batch_decode.py ...

Prints one JSON line (rank 0): aggregate decode tok/s (device-timed, max over ranks), per-step ms, and the fraction of the HBM
roofline (weights 0.858 GB per step shared by the B rows + each row's KV read)."""
import argparse
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from mllm_b200 import mllm_file as mf  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=8)
    ap.add_argument("--prompt", type=int, default=32)
    ap.add_argument("--decode", type=int, default=200)
    ap.add_argument("--model", default="qwen1.5-1.8b")
    ap.add_argument("--reps", type=int, default=3)
    a = ap.parse_args()
    import torch
    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    from mllm_b200 import backend
    cfg = mf.config(a.model, cache_limit=a.prompt + a.decode + 8)
    path = f"/tmp/{a.model}-q4_0-r{local}.mllm"
    bench.make_model_file(cfg, path, device=f"cuda:{local}")
    model = backend.LlamaModel(cfg, device=local).load(path)
    rng = np.random.default_rng(1234 + rank)  # every rank serves its own requests
    ids = rng.integers(0, cfg["vocab"], (a.batch, a.prompt)).astype(np.float32)
    best = None
    for rep in range(a.reps + 1):
        model.clear_kvcache()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        ms_p, ms_d, toks = model.request_device(ids, a.decode)
        t = torch.tensor([ms_p, ms_d], dtype=torch.float64, device="cuda")
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        if rep > 0 and (best is None or t[1].item() < best[1]):
            best = (t[0].item(), t[1].item())
    fused = model.decode_mode()
    if rank == 0:
        ms_p, ms_d = best
        step_ms = ms_d / a.decode
        wbytes = mf.decode_weight_bytes(cfg) if hasattr(mf, "decode_weight_bytes") else None
        hd = cfg["hidden"] // cfg["heads"]
        ctx = a.prompt + a.decode / 2
        kv = 2 * cfg["layers"] * ctx * cfg["kv_heads"] * hd * 2 * a.batch
        peak = bench.measured_peak_gbs() if hasattr(bench, "measured_peak_gbs") else None
        out = {"metric": f"{a.model}_q4_0_batch{a.batch}_decode_tok_s", "value": world * a.batch * a.decode / (ms_d * 1e-3), "unit": "tok/s", "n_gpus": world,
               "batch_per_gpu": a.batch, "prompt": a.prompt, "decode": a.decode, "ms_per_step": step_ms, "prefill_ms": ms_p, "decode_mode": fused,
               "scaling": "weak", "parallelism": f"dp{world} replicas, requests sharded, no collective"}
        if wbytes:
            gb = (wbytes + kv) / 1e9
            out["algorithmic_GB_per_step"] = gb
            out["GBs"] = gb / (step_ms * 1e-3)
            if peak:
                out["frac_of_hbm_peak"] = out["GBs"] / peak
        print(json.dumps(out))
    model.close()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
