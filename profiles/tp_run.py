"""Tensor-parallel run of a LLaMA-family Q4_0 model (mllm_b200/tp.py): one process per GPU.
  python profiles/tp_run.py --model toy-gqa --prompt 16 --decode 24 [--save out.npz]                      (1 GPU, no exchange)
  python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29650 profiles/tp_run.py --model llama2-7b ...
Rank 0 prints one JSON line: tokens, prefill ms, decode ms/token (CUDA events, max over ranks), launches per step, exchange path."""
import argparse
import json
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from mllm_b200 import build, mllm_file as mf, tp  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--model", default="llama2-7b")
    ap.add_argument("--prompt", type=int, default=512)
    ap.add_argument("--decode", type=int, default=64)
    ap.add_argument("--limit", type=int, default=1024)
    ap.add_argument("--seed", type=int, default=1234)
    ap.add_argument("--save", default="")
    ap.add_argument("--no-graph", action="store_true")
    a = ap.parse_args()
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    if rank == 0:
        build.build_all()
    if dist:
        dist.barrier()
    cfg = mf.config(a.model, cache_limit=a.limit)
    path = f"/tmp/mllm_tp_{a.model}_q4_0.mllm"
    if rank == 0:
        if a.model.startswith("toy"):
            mf.write_model(path, cfg, mf.random_tensors(cfg, seed=a.seed))
        else:
            mf.write_random_model_torch(cfg, path, device="cuda", seed=a.seed)
    if dist:
        dist.barrier()
    m = tp.TPLlama(cfg, path, rank, world, dist)
    ids = np.random.default_rng(a.seed).integers(0, cfg["vocab"], size=(1, a.prompt)).astype(np.float32)
    e = [torch.cuda.Event(enable_timing=True) for _ in range(4)]
    for rep in range(3):  # the last request is the timed one (graph built, scratch allocated, NCCL warmed up)
        m.clear()
        torch.cuda.synchronize()
        if dist:
            dist.barrier()
        e[0].record()
        logits0 = m.prefill(ids)
        first = logits0.argmax(dim=1)
        e[1].record()
        torch.cuda.synchronize()
        if dist:
            dist.barrier()
        e[2].record()
        toks, logits = m.decode(first.cpu().numpy(), a.decode, graph=not a.no_graph)
        e[3].record()
        torch.cuda.synchronize()
    t = torch.tensor([e[0].elapsed_time(e[1]), e[2].elapsed_time(e[3])], device="cuda")
    if dist:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    tokens = [int(first[0])] + [int(x[0]) for x in toks]
    if a.save and rank == 0:
        np.savez(a.save, logits0=logits0.cpu().numpy(), logits=logits.cpu().numpy(), tokens=np.array(tokens))
    if dist:  # every rank must have picked the same tokens
        allt = [None] * world
        dist.all_gather_object(allt, tokens)
        assert all(x == allt[0] for x in allt), "ranks disagree on the tokens"
    if rank == 0:
        print(json.dumps({"model": a.model, "tp": world, "exchange": "none" if world == 1 else (("p2p-fused-in-gemv" if m.fused_exchange else "p2p-kernel") if m.use_p2p else "nccl"), "prompt": a.prompt, "decode": a.decode,
                          "prefill_ms": round(float(t[0]), 3), "decode_ms_per_token": round(float(t[1]) / a.decode, 4), "launches_per_step": m.launches_per_step,
                          "graph": not a.no_graph, "tokens": tokens[:32]}), flush=True)
    m.close()
    if dist:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
