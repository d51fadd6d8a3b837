"""N > 1 logic of bench.py on CPU: two gloo ranks, max-over-ranks timing, whole-job aggregation, and the reference arm
running on rank 0 only."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

WORKER = r"""
import os, sys, json
sys.path.insert(0, os.environ["REPO"])
import torch.distributed as dist
import bench
dist.init_process_group("gloo")
rank, world = dist.get_rank(), dist.get_world_size()
mine = [100.0 + 50.0 * rank, 7.0 - rank]            # rank 1 is slower on the first, faster on the second
got = bench.max_over_ranks(mine, world, "cpu")
dist.barrier()
if rank == 0:
    print(json.dumps({"max": got, "value": bench.job_tokens_per_s(768, world, got[0])}))
dist.destroy_process_group()
"""


def test_two_ranks_max_time_and_aggregate(tmp_path):
    w = tmp_path / "worker.py"
    w.write_text(WORKER)
    env = dict(os.environ, REPO=ROOT)
    out = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
                          "--master-port", "29631", str(w)], env=env, capture_output=True, text=True, timeout=300)
    assert out.returncode == 0, out.stderr[-2000:]
    line = [l for l in out.stdout.splitlines() if l.startswith("{")][-1]
    r = json.loads(line)
    assert r["max"] == [150.0, 7.0]
    assert abs(r["value"] - 768 * 2 / 150.0 * 1e3) < 1e-6


def test_single_rank_passthrough():
    sys.path.insert(0, ROOT)
    import bench
    assert bench.max_over_ranks([3, 4.5], 1, "cpu") == [3.0, 4.5]
    assert bench.job_tokens_per_s(768, 1, 768.0) == 1000.0


def test_reference_arm_runs_on_rank0_only():
    sys.path.insert(0, ROOT)
    import bench

    class A:
        warmup, steps, cpu_prefill, cpu_decode = 0, 1, 1, 0
    assert bench.reference_arm(A, rank=1, world=2) is None  # other ranks exit without work
