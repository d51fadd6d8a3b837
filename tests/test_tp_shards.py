"""Host logic of the optional tensor-parallel Linear (mllm_b200/tp.py), on CPU: the shard boundaries, the slicing of Q4_0 parameters
(column-parallel rows, row-parallel whole-block K slices), the rotary tables against the oracle's, and -- two gloo ranks -- that each
rank's slice of a layer, run through the oracle's kernels (the checker; tests may), reduces to the unsharded layer."""
import json
import os
import subprocess
import sys

import numpy as np
import pytest

from mllm_b200 import mllm_file as mf, tp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_split_units_covers_and_balances():
    for total, unit, world in ((11008, 128, 8), (11008, 128, 4), (11008, 128, 2), (512, 128, 2), (5504, 128, 8), (4096, 128, 1)):
        parts = tp.split_units(total, unit, world)
        assert parts[0][0] == 0 and parts[-1][1] == total
        assert all(a[1] == b[0] for a, b in zip(parts, parts[1:]))
        sizes = [hi - lo for lo, hi in parts]
        assert all(s % unit == 0 and s > 0 for s in sizes) and max(sizes) - min(sizes) <= unit
    assert [hi - lo for lo, hi in tp.split_units(11008, 128, 8)] == [1408] * 6 + [1280] * 2
    with pytest.raises(ValueError):
        tp.split_units(100, 128, 2)
    with pytest.raises(ValueError):
        tp.split_units(128, 128, 2)


def test_shard_plan_llama2_7b():
    cfg = mf.config("llama2-7b")
    for world in (1, 2, 4, 8):
        plan = tp.shard_plan(cfg, world)
        assert [p["q"][1] - p["q"][0] for p in plan] == [32 // world] * world
        assert plan[-1]["ffn"][1] == 11008
    with pytest.raises(ValueError):
        tp.shard_plan(mf.config("toy-gqa"), 4)  # 2 kv heads do not divide over 4 ranks


def test_shards_reassemble_to_the_parameters():
    cfg = mf.config("toy-gqa")
    t = mf.random_tensors(cfg, seed=3)
    n = mf.Names(cfg["rope_type"])
    H, D, F = cfg["hidden"], cfg["head_dim"], cfg["ffn"]
    sh = [tp.shard_tensors(t, cfg, r, 2) for r in range(2)]
    for role in ("wq", "wk", "wv", "w_gate", "w_up"):
        full = np.asarray(t[n.layer(1, role)]).reshape(-1, H // 32 * 18)
        assert np.array_equal(np.concatenate([s[n.layer(1, role)] for s in sh]), full)
    for role, K in (("wo", cfg["heads"] * D), ("w_down", F)):
        full = np.asarray(t[n.layer(0, role)]).reshape(H, K // 32, 18)
        got = np.concatenate([s[n.layer(0, role)].reshape(H, -1, 18) for s in sh], axis=1)
        assert np.array_equal(got, full)
    assert np.array_equal(sh[0][n.embed].reshape(-1), np.asarray(t[n.embed]).reshape(-1)) and sh[0][n.embed].shape[0] == cfg["vocab"] and np.array_equal(sh[1][n.norm], t[n.norm])


@pytest.mark.parametrize("rt", [2, 4])
def test_rope_tables_equal_the_oracles(oracle, rt):
    s, c = tp.rope_tables(rt, 40, 64, 10000.0)
    so, co = oracle.only.rope_tables(rt, 40, 64, 10000.0)
    assert np.array_equal(s.view(np.uint32), so.view(np.uint32)) and np.array_equal(c.view(np.uint32), co.view(np.uint32))


WORKER = r"""
import os, sys, json
sys.path.insert(0, os.environ["REPO"])
import numpy as np
import torch, torch.distributed as dist
from mllm_b200 import mllm_file as mf, tp
from oracle import pyoracle
dist.init_process_group("gloo")
rank, world = dist.get_rank(), dist.get_world_size()
orc = pyoracle.port()
cfg = mf.config("toy-gqa")
t = mf.random_tensors(cfg, seed=3)
n = mf.Names(cfg["rope_type"])
sh = tp.shard_tensors(t, cfg, rank, world)
H, F = cfg["hidden"], cfg["ffn"]
rng = np.random.default_rng(0)
x = rng.standard_normal((5, H)).astype(np.float32)
# FFN of layer 0: column-parallel gate / up on the slice, row-parallel down, all-reduce of the partial sums
f0, f1 = tp.shard_plan(cfg, world)[rank]["ffn"]
g = orc.linear(sh[n.layer(0, "w_gate")], x)
u = orc.linear(sh[n.layer(0, "w_up")], x)
a = orc.silu(g) * u
part = torch.from_numpy(orc.linear(sh[n.layer(0, "w_down")], a))
dist.all_reduce(part)
# unsharded
G = orc.linear(np.asarray(t[n.layer(0, "w_gate")]).reshape(F, -1), x)
U = orc.linear(np.asarray(t[n.layer(0, "w_up")]).reshape(F, -1), x)
A = orc.silu(G) * U
full = orc.linear(np.asarray(t[n.layer(0, "w_down")]).reshape(H, -1), A)
ok_cols = bool(np.array_equal(g, G[:, f0:f1]) and np.array_equal(u, U[:, f0:f1]))
err = float(np.abs(part.numpy() - full).max() / np.abs(full).max())
res = [None] * world
dist.all_gather_object(res, {"rank": rank, "cols_exact": ok_cols, "rel_err": err})
if rank == 0:
    print(json.dumps(res))
dist.destroy_process_group()
"""


def test_two_gloo_ranks_reduce_to_the_unsharded_ffn(tmp_path, oracle):
    w = tmp_path / "worker.py"
    w.write_text(WORKER)
    env = dict(os.environ, REPO=ROOT, OMP_NUM_THREADS="1")
    out = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
                          "--master-port", "29641", str(w)], env=env, capture_output=True, text=True, timeout=300)
    assert out.returncode == 0, out.stderr[-2000:]
    res = json.loads([l for l in out.stdout.splitlines() if l.startswith("[")][-1])
    assert len(res) == 2 and all(r["cols_exact"] for r in res)
    # the partial sums meet in another order than the single dot product's, and the activation slices are re-quantised per slice
    # with the SAME per-32 blocks (cuts fall on block boundaries), so only the fp32 summation order differs
    assert all(r["rel_err"] < 1e-5 for r in res), res
