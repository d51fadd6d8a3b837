#!/usr/bin/env python
"""bench.py -- LLaMA-2-7B Q4_0, prefill 512 + decode 256 (BASELINE.json configs[1]) on the CUDA backend.

A step = one request: clear the KV cache, prefill 512 synthetic tokens, then 256 greedy decode steps.
  value : tokens/s of the whole request with inputs resident in HBM (device-timed, CUDA events).
  e2e   : the same request through the host-buffer API a user calls (LlamaModel.generate: prompt ids H2D,
          prefill + greedy decode with the pick made on the device, the 256 token ids D2H), wall clock
          around the synchronous call.
  roofline : the decode GEMV (k_gemv*, the dominant kernel) -- algorithmic weight bytes / launch time,
          measured live with CUDA events on the backend stream over all Linear shapes of a decode step.
  cpu_baseline : the reference's CPU kernels (oracle/_ref) or the port, on a bounded sample.
`--impl reference` times only the CPU arm.  N>1: one replica per GPU, no collective ("weak").
"""
import argparse
import faulthandler
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
faulthandler.enable()  # a native crash leaves a Python traceback on stderr instead of a silent exit

from mllm_b200 import mllm_file as mf  # noqa: E402

PREFILL, DECODE, LIMIT = 512, 256, 1024
MODEL = "llama2-7b"
METRIC = "llama2_7b_q4_0_prefill512_decode256_tok_s"


def log(*a):
    print(*a, file=sys.stderr, flush=True)


# ----------------------------------------------------------------------------- model file
def make_model_file(cfg, path, device="cuda", seed=1234):
    """Random-init .mllm (W ~ N(0,0.02^2) -> Q4_0 with the reference quantiser's arithmetic, norms 1+N(0,0.02^2))."""
    return mf.write_random_model_torch(cfg, path, device=device, seed=seed)


# ----------------------------------------------------------------------------- clocks
class ClockSampler:
    Q = "clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index):
        self.index, self.rows, self.p = index, [], None

    def start(self):
        try:
            self.p = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "200"],
                                      stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.p = None

    def _read(self):
        for line in self.p.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if self.p is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.p.terminate()
        try:
            self.p.wait(timeout=3)
        except Exception:
            self.p.kill()
        sm, mx, reasons = [], None, set()
        for r in self.rows:
            try:
                sm.append(float(r[0]))
                mx = float(r[1])
                for name, v in zip(["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"], r[3:7]):
                    if v.lower().startswith("active"):
                        reasons.add(name)
            except Exception:
                pass
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": mx, "reasons": sorted(reasons), "samples": len(sm)}


# ----------------------------------------------------------------------------- CPU arm
def _host_cpus():
    try:
        return len(os.sched_getaffinity(0))
    except Exception:
        return os.cpu_count() or 1


def _cpu_runner(cfg, path, threads, limit):
    """The reference's own model graph on its CPU backend (oracle/_ref/libmllm_cpu_ref.so: LLaMAModel on MLLM_CPU, compiled from the
    reference's sources) -- else the reference's kernels under the C graph (libmllm_ref.so) -- else the C port."""
    from oracle import pyoracle as po
    c = dict(cfg, cache_limit=limit)
    if po.cpuref() is not None:
        return po.RefGraph(c, path, threads=threads), "reference", "reference LLaMAModel graph on CPUBackend (oracle/_ref/libmllm_cpu_ref.so)"
    po.build()
    L, kind = po.ref(), "reference"
    if L is None:
        L, kind = po.port(), "port"
    os.environ["OMP_NUM_THREADS"] = str(threads)
    tensors, _ = mf.read_mllm(path)
    return po.ModelRunner(L, cfg, tensors, mf.Names(cfg["rope_type"]), cache_limit=limit), kind, "reference ggml/FA2 kernels under oracle/model_runner.c"


def cpu_arm(cfg, path, prefill, decode, reps=1):
    """Times the CPU reference on a bounded sample of the 512 + 256 request: `prefill` prompt tokens, then `decode` greedy steps.
    Thread count is chosen explicitly (never inherited: torchrun exports OMP_NUM_THREADS=1): all host CPUs, or half of them when
    that decodes faster (SMT siblings / oversubscribed vCPUs stall the reference's spin-waiting OpenMP teams)."""
    n = _host_cpus()
    rng = np.random.default_rng(7)
    ids = rng.integers(0, cfg["vocab"], (1, prefill)).astype(np.float32)
    probe = {}
    for th in sorted({n, max(1, n // 2)}, reverse=True):
        r, kind, how = _cpu_runner(cfg, path, th, prefill + decode + 8)
        lg = r.forward(ids[:, :4])
        t0 = time.perf_counter()
        for _ in range(2):
            lg = r.forward(np.array([[lg.argmax()]], dtype=np.float32))
        probe[th] = (time.perf_counter() - t0) / 2
        if hasattr(r, "close"):
            r.close()
    threads = min(probe, key=probe.get)
    runner, kind, how = _cpu_runner(cfg, path, threads, prefill + decode + 8)
    times = []
    for _ in range(reps):
        runner.reset()
        t0 = time.perf_counter()
        lg = runner.forward(ids)
        t1 = time.perf_counter()
        for _ in range(decode):
            lg = runner.forward(np.array([[lg.argmax()]], dtype=np.float32))
        t2 = time.perf_counter()
        times.append((t1 - t0, t2 - t1))
    tp = float(np.mean([t[0] for t in times]))
    td = float(np.mean([t[1] for t in times]))
    pre_tok_s, dec_tok_s = prefill / tp, (decode / td if decode else None)
    # the metric's own request (512 + 256) at the sampled per-phase rates
    projected = (PREFILL + DECODE) / (PREFILL / pre_tok_s + DECODE / dec_tok_s) if dec_tok_s else pre_tok_s
    return {"value": projected, "unit": "tok/s", "cores": threads, "host_cpus": n, "kind": kind,
            "sample": f"{how}; full {cfg['name']} Q4_0 model, prefill {prefill} + decode {decode} tokens timed on {threads} threads "
                      f"(decode probe s/token by thread count: { {k: round(v, 4) for k, v in probe.items()} }), value = the 512 + 256 request at the sampled per-phase rates",
            "prefill_tok_s": pre_tok_s, "decode_tok_s": dec_tok_s, "seconds": tp + td}


# ----------------------------------------------------------------------------- ranks
def max_over_ranks(values, world, device):
    """Each rank times its own replica; the job's time is the slowest rank's (no collective on the data path)."""
    if world == 1:
        return [float(v) for v in values]
    import torch
    import torch.distributed as dist
    t = torch.tensor(values, device=device, dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return [float(v) for v in t.tolist()]


def job_tokens_per_s(tokens_per_replica, world, ms):
    """Whole-job throughput of `world` independent replicas that each processed tokens_per_replica in <= ms."""
    return tokens_per_replica * world / ms * 1e3


# ----------------------------------------------------------------------------- GPU arm
def gemv_roofline(cfg, peaks, M=1):
    """The dominant kernel of the request is the decode Linear, k_gemv_tma (fused activation quantise + Q4_0 GEMV).
    Every Linear launch of one decode step (qkv, o, gate|up, down per block + lm_head) is timed live: the launches are
    captured in a CUDA graph (device time, not ctypes dispatch), weights rotate over buffers larger than L2 (cold HBM
    reads), CUDA events on the stream the graph is launched on.  achieved = algorithmic weight bytes / launch time,
    launch-weighted over the step."""
    import torch
    from mllm_b200 import capi
    H, F, V, D = cfg["hidden"], cfg["ffn"], cfg["vocab"], cfg["head_dim"]
    kvd = cfg["kv_heads"] * D
    L = cfg["layers"]
    shapes = [("qkv", [H, kvd, kvd], H, 0, True, L), ("o", [H], H, 1, False, L), ("gate_up", [F, F], H, 2, True, L),
              ("down", [H], F, 1, False, L), ("lm_head", [V], H, 0, True, 1)]
    total_t, total_b, per = 0.0, 0.0, {}
    for tag, Ns, K, epi, norm, count in shapes:
        wbytes = sum(Ns) * (K // 32) * 18
        nbuf = max(2, int(200e6 // wbytes) + 1)
        bufs = [[torch.randint(0, 255, (n, K // 32 * 18), dtype=torch.uint8, device="cuda") for n in Ns] for _ in range(nbuf)]
        x = torch.randn((M, K), device="cuda")
        nw = torch.ones(K, device="cuda") if norm else None
        res = torch.randn((M, Ns[0]), device="cuda") if epi == 1 else None

        def call(ws):
            capi.gemv_fused(ws, x, norm_w=nw, epilogue=epi, residual=res, pdl=True)
        for i in range(3):
            call(bufs[i % nbuf])
        reps = 4 * nbuf
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g, stream=side):
                for i in range(reps):
                    call(bufs[i % nbuf])
            g.replay()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            torch.cuda.synchronize()
            e0.record()
            g.replay()
            e1.record()
            torch.cuda.synchronize()
        t = e0.elapsed_time(e1) / reps * 1e-3
        per[tag] = {"N": sum(Ns) if epi != 2 else Ns[0], "K": K, "us": round(t * 1e6, 2), "GBs": round(wbytes / t / 1e9, 1), "launches_per_step": count,
                    "bytes": wbytes}
        total_t += t * count
        total_b += wbytes * count
        del bufs, g
    achieved = total_b / total_t / 1e9
    peak = peaks.get("hbm_gbs", 6650.0)
    return {"bound": "hbm", "achieved": round(achieved, 1), "peak": peak, "unit": "GB/s", "frac": round(achieved / peak, 4),
            # DRAM bytes of one gate|up launch (its largest share of the step) from the ncu --set full capture summarised in
            # profiles/r01_ncu_gemv_gateup.txt: 50.80 MB read + 3.6 KB written against 50.72 MB algorithmic -- no re-reads
            "traffic": 50802688, "traffic_source": "profiles/r01_ncu_gemv_gateup.txt (one ncu --set full capture, NOT measured by this run)",
            "traffic_of": "one gate|up launch (k_gemv_tma<1,1,2,1>): dram read 50.799 MB + write 0.004 MB vs 50.725 MB algorithmic",
            "kernel": ("k_gemv_tma (decode Linear: fused RMSNorm/Q8_0 prologue + Q4_0 GEMV)" if M < 3 else
                       f"k_gemv_mma (batched decode Linear, {M} rows: mma.sync.m16n8k32.u8.s8 over TMA-staged Q4_0 tiles) + k_act_quant_mma") + ", launch-weighted over one decode step",
            "peak_source": "measured (MEASURED_PEAKS.json hbm_gbs)" if "hbm_gbs" in peaks else "fallback 6650",
            "per_shape": per}


def gpu_arm(args, rank, world, local_rank):
    import torch
    torch.cuda.set_device(local_rank)
    from mllm_b200 import backend, build
    if rank == 0:
        build.build_all()  # a no-op when the in-tree libraries are current; never two ranks compiling the same objects
    if world > 1:
        import torch.distributed as dist
        dist.barrier()
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    cfg = mf.config(MODEL, cache_limit=LIMIT)
    path = os.environ.get("MLLM_BENCH_MODEL", f"/tmp/mllm_bench_{MODEL}_q4_0.mllm")
    t0 = time.time()
    if rank == 0:
        make_model_file(cfg, path, device=f"cuda:{local_rank}")
    if world > 1:
        import torch.distributed as dist
        dist.barrier()
    log(f"[rank {rank}] model file ready in {time.time() - t0:.1f}s")
    t0 = time.time()
    model = backend.LlamaModel(cfg, device=local_rank).load(path)
    log(f"[rank {rank}] loaded {mf.model_bytes(cfg) / 1e9:.2f} GB through ParamLoader in {time.time() - t0:.1f}s")
    rng = np.random.default_rng(1234 + rank)
    ids = rng.integers(0, cfg["vocab"], (1, PREFILL)).astype(np.float32)

    def request():
        model.clear_kvcache()
        return model.request_device(ids, DECODE)

    for _ in range(args.warmup):
        request()
    sampler = ClockSampler(local_rank)
    sampler.start()
    if world > 1:
        import torch.distributed as dist
        dist.barrier()
    torch.cuda.synchronize()
    l0 = backend.kernel_launches()
    tp, td = [], []
    for _ in range(args.steps):
        a, b, _ = request()
        tp.append(a)
        td.append(b)
    torch.cuda.synchronize()
    launches = backend.kernel_launches() - l0
    total_ms = float(np.sum(tp) + np.sum(td))
    # e2e: host-buffer API (prompt ids H2D, device-side greedy loop, token ids D2H), wall clock around the synchronous call
    e2e_ms = []
    for _ in range(args.steps):
        model.clear_kvcache()
        w0 = time.perf_counter()
        model.generate(ids, DECODE)
        e2e_ms.append((time.perf_counter() - w0) * 1e3)
    clocks = sampler.stop()
    total_ms, e2e_mean = max_over_ranks([total_ms, float(np.mean(e2e_ms))], world, "cuda")
    if rank != 0:
        return
    ms_per_step = total_ms / args.steps
    toks = (PREFILL + DECODE) * world  # == job_tokens_per_s numerator
    roof = gemv_roofline(cfg, peaks)
    wbytes = mf.decode_weight_bytes(cfg)
    head_bytes = mf.nbytes(cfg["vocab"], cfg["hidden"], mf.Q4_0)
    kv_bytes = 2 * cfg["layers"] * (PREFILL + DECODE / 2) * cfg["kv_heads"] * cfg["head_dim"] * 2
    dec_ms = float(np.mean(td)) / DECODE
    out = {
        "metric": METRIC, "value": round(job_tokens_per_s(PREFILL + DECODE, world, ms_per_step), 2), "unit": "tok/s", "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": round(ms_per_step, 3), "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "int8 block dot (Q4_0 x Q8_0) + fp32 accumulate, fp16 KV", "data": "synthetic",
        "config": {"workload": "LLaMA-2-7B Q4_0 (BASELINE configs[1]): prefill 512 + greedy decode 256, batch 1, cache limit 1024, "
                               "random-init .mllm loaded through ParamLoader; one replica per GPU",
                   "l2": "weights 3.7 GB per step >> 126 MB L2 (inputs larger than L2)", "parallelism": f"dp{world} replicas, no collective"},
        "prefill_tok_s": round(PREFILL / (float(np.mean(tp)) * 1e-3), 1), "decode_tok_s": round(1e3 / dec_ms, 1),
        "decode_step": {"ms": round(dec_ms, 4), "algorithmic_GB": round((wbytes + kv_bytes) / 1e9, 4),
                        "GBs": round((wbytes + kv_bytes) / (dec_ms * 1e-3) / 1e9, 1),
                        "frac_of_hbm_peak": round((wbytes + kv_bytes) / (dec_ms * 1e-3) / 1e9 / peaks.get("hbm_gbs", 6650.0), 4)},
        "prefill": {"ms": round(float(np.mean(tp)), 3), "int8_TOPs": round(2 * ((wbytes - head_bytes) / 0.5625) * PREFILL / (float(np.mean(tp)) * 1e-3) / 1e12, 1),
                    "frac_of_int8_peak": round(2 * ((wbytes - head_bytes) / 0.5625) * PREFILL / (float(np.mean(tp)) * 1e-3) / 1e12 / (2 * peaks.get("bf16_tflops", 1656.9)), 4),
                    "int8_peak_TOPs": round(2 * peaks.get("bf16_tflops", 1656.9), 1), "note": "Linear MACs x 2 of the 512 prompt tokens (lm_head runs on the last token only and is not counted); peak = 2 x measured dense bf16"},
        "e2e": {"value": round(job_tokens_per_s(PREFILL + DECODE, world, e2e_mean), 2), "unit": "tok/s", "h2d_bytes_per_step": PREFILL * 4,
                # token ids of the first pick (fp32 row per step) + the device-side token log the fused steps write (one int32 per position)
                "d2h_bytes_per_step": DECODE * 4 + (LIMIT + 1) * 4, "ms_per_step": round(e2e_mean, 2)},
        "decode_mode": model.decode_mode(), "gpu_launches": int(launches), "clocks": clocks, "roofline": roof,
    }
    if not args.no_cpu:
        # the CPU arm is a reported baseline: it runs in its own process so that nothing it does (OpenMP runtime, a crash,
        # a stall) can take the GPU line down
        try:
            r = subprocess.run([sys.executable, os.path.abspath(__file__), "--impl", "reference", "--steps", "1", "--warmup", "0",
                                "--cpu-prefill", str(args.cpu_prefill), "--cpu-decode", str(args.cpu_decode)],
                               capture_output=True, text=True, timeout=900,
                               env={k: v for k, v in dict(os.environ, RANK="0", WORLD_SIZE="1", LOCAL_RANK="0").items() if not k.startswith("OMP_")})
            line = [l for l in r.stdout.splitlines() if l.startswith("{")][-1]
            out["cpu_baseline"] = json.loads(line)["cpu_baseline"]
        except Exception as e:
            out["cpu_baseline"] = {"error": repr(e)[:300]}
    print(json.dumps(out), flush=True)


def gpu_arm_c5(args, rank, world, local_rank):
    """BASELINE configs[4]: Qwen1.5-1.8B Q4_0 batched greedy decode, the demo_qwen_batch shape (examples/demo_qwen_batch.cpp:61-68:
    Module::generate over a batch of prompts, 200 new tokens) with B requests of 32 prompt tokens PER GPU; requests are sharded
    data-parallel, one replica per GPU, no collective (weak scaling).  A step = one batch request (prefill + 200 greedy steps);
    value = decode tokens/s of the whole job (B x 200 x N / slowest rank's decode time), device-timed."""
    import torch
    torch.cuda.set_device(local_rank)
    from mllm_b200 import backend, build
    if rank == 0:
        build.build_all()
    if world > 1:
        import torch.distributed as dist
        dist.barrier()
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    B, S, ND = args.batch, 32, 200
    cfg = mf.config("qwen1.5-1.8b", cache_limit=S + ND + 8)
    path = os.environ.get("MLLM_BENCH_MODEL_C5", "/tmp/mllm_bench_qwen1.5-1.8b_q4_0.mllm")
    if rank == 0:
        make_model_file(cfg, path, device=f"cuda:{local_rank}")
    if world > 1:
        import torch.distributed as dist
        dist.barrier()
    model = backend.LlamaModel(cfg, device=local_rank).load(path)
    ids = np.random.default_rng(1234 + rank).integers(0, cfg["vocab"], (B, S)).astype(np.float32)  # every rank serves its own requests

    def request():
        model.clear_kvcache()
        return model.request_device(ids, ND)

    for _ in range(args.warmup):
        request()
    sampler = ClockSampler(local_rank)
    sampler.start()
    if world > 1:
        import torch.distributed as dist
        dist.barrier()
    torch.cuda.synchronize()
    l0 = backend.kernel_launches()
    tp, td = [], []
    for _ in range(args.steps):
        a, b, _ = request()
        tp.append(a)
        td.append(b)
    torch.cuda.synchronize()
    launches = backend.kernel_launches() - l0
    e2e_ms = []
    for _ in range(args.steps):
        model.clear_kvcache()
        w0 = time.perf_counter()
        model.generate(ids, ND)
        e2e_ms.append((time.perf_counter() - w0) * 1e3)
    clocks = sampler.stop()
    dec_ms, pre_ms, e2e_mean = max_over_ranks([float(np.sum(td)), float(np.sum(tp)), float(np.mean(e2e_ms))], world, "cuda")
    if rank != 0:
        return
    step_ms = dec_ms / args.steps / (ND - 1)
    wbytes = mf.decode_weight_bytes(cfg)
    kv = 2 * cfg["layers"] * (S + ND / 2) * cfg["kv_heads"] * cfg["head_dim"] * 2 * B
    peak = peaks.get("hbm_gbs", 6650.0)
    roof = gemv_roofline(cfg, peaks, M=B)
    roof["traffic"] = None
    roof.pop("traffic_source", None)
    roof.pop("traffic_of", None)
    out = {
        "metric": f"qwen1.5_1.8b_q4_0_batch{B}_decode_tok_s", "value": round(B * (ND - 1) * world / (dec_ms / args.steps) * 1e3, 1), "unit": "tok/s",
        "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": round((dec_ms + pre_ms) / args.steps, 3), "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "int8 block dot (Q4_0 x Q8_0) + fp32 accumulate, fp16 KV", "data": "synthetic",
        "config": {"workload": f"Qwen1.5-1.8B Q4_0 (BASELINE configs[4], demo_qwen_batch shape): {B} requests per GPU, prompt {S} + {ND} greedy tokens, "
                               "random-init .mllm through ParamLoader; requests sharded data-parallel, one replica per GPU",
                   "l2": "weights 0.86 GB per decode step >> 126 MB L2 (inputs larger than L2)", "parallelism": f"dp{world} replicas, no collective"},
        "decode_step": {"ms": round(step_ms, 4), "algorithmic_GB": round((wbytes + kv) / 1e9, 4), "GBs": round((wbytes + kv) / (step_ms * 1e-3) / 1e9, 1),
                        "frac_of_hbm_peak": round((wbytes + kv) / (step_ms * 1e-3) / 1e9 / peak, 4)},
        "prefill": {"ms": round(pre_ms / args.steps, 3), "tokens": B * S},
        "e2e": {"value": round(B * ND * world / e2e_mean * 1e3, 1), "unit": "tok/s (prefill + decode wall clock, all generated tokens)", "h2d_bytes_per_step": B * S * 4,
                "d2h_bytes_per_step": B * ND * 4 + (S + ND + 9) * B * 4, "ms_per_step": round(e2e_mean, 2)},
        "decode_mode": model.decode_mode(), "gpu_launches": int(launches), "clocks": clocks, "roofline": roof,
    }
    if not args.no_cpu:
        try:  # the reference graph (QWenForCausalLM on MLLM_CPU) on a bounded sample: 3 requests (the demo's batch), 32 + 8 tokens
            from oracle import pyoracle as po
            n = _host_cpus()
            th = max(1, n // 2)
            g = po.RefGraph(dict(cfg, cache_limit=64), path, threads=th)
            cid = ids[:3]
            t0 = time.perf_counter()
            g.generate(cid, 8)
            dt = time.perf_counter() - t0
            g.close()
            out["cpu_baseline"] = {"value": round(3 * 8 / dt, 2), "unit": "tok/s", "cores": th, "host_cpus": n, "kind": "reference",
                                   "sample": "reference QWenForCausalLM graph on CPUBackend (oracle/_ref/libmllm_cpu_ref.so), Module::generate on 3 requests x (32 prompt + 8 new tokens), wall clock incl. prefill"}
        except Exception as e:
            out["cpu_baseline"] = {"error": repr(e)[:300]}
    print(json.dumps(out), flush=True)


def gpu_arm_tp(args, rank, world, local_rank):
    """The optional tensor-parallel path (north_star): ONE LLaMA-2-7B Q4_0 request (BASELINE configs[1]: prefill 512 + 256
    greedy tokens) with every Linear sharded over the N ranks (mllm_b200/tp.py: column-parallel q/k/v/gate/up, row-parallel o/down,
    NCCL all-reduce in prefill, the peer-memory exchange kernel of csrc/p2p.cu per decode token; MLLM_TP_NCCL=1: NCCL there too).
    Strong scaling: the work is fixed, value = the request's tokens / slowest rank's device time."""
    import torch
    torch.cuda.set_device(local_rank)
    from mllm_b200 import build, tp
    dist = None
    if world > 1:
        import torch.distributed as dist
    if rank == 0:
        build.build_all()
    if dist:
        dist.barrier()
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    S, ND = 512, 256
    cfg = mf.config("llama2-7b", cache_limit=1024)
    path = os.environ.get("MLLM_BENCH_MODEL", "/tmp/mllm_bench_llama2-7b_q4_0.mllm")
    if rank == 0:
        make_model_file(cfg, path, device=f"cuda:{local_rank}")
    if dist:
        dist.barrier()
    model = tp.TPLlama(cfg, path, rank, world, dist)
    ids = np.random.default_rng(1234).integers(0, cfg["vocab"], (1, S)).astype(np.float32)
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(3)]

    def request():
        model.clear()
        ev[0].record()
        first = model.prefill(ids).argmax(dim=1)
        ev[1].record()
        toks, _ = model.decode(first.cpu().numpy(), ND - 1)
        ev[2].record()
        torch.cuda.synchronize()
        return ev[0].elapsed_time(ev[1]), ev[1].elapsed_time(ev[2]), [int(first[0])] + [int(t[0]) for t in toks]

    for _ in range(max(args.warmup, 1)):
        request()
    sampler = ClockSampler(local_rank)
    sampler.start()
    if dist:
        dist.barrier()
    tp_ms, td_ms, e2e = [], [], []
    for _ in range(args.steps):
        w0 = time.perf_counter()
        a, b, toks = request()
        e2e.append((time.perf_counter() - w0) * 1e3)
        tp_ms.append(a)
        td_ms.append(b)
    clocks = sampler.stop()
    pre, dec, wall = max_over_ranks([float(np.mean(tp_ms)), float(np.mean(td_ms)), float(np.mean(e2e))], world, "cuda")
    launches = model.launches_per_step * (ND - 1) * args.steps
    exchange = "none" if world == 1 else (("p2p-fused-in-gemv" if model.fused_exchange else "p2p-kernel") if model.use_p2p else "nccl")
    model.close()
    if rank != 0:
        return
    step_ms = dec / (ND - 1)
    wbytes = mf.decode_weight_bytes(cfg)
    lm = cfg["vocab"] * cfg["hidden"] // 32 * 18  # the lm_head stream is replicated on every rank
    per_rank = (wbytes - lm) / world + lm
    peak = peaks.get("hbm_gbs", 6650.0)
    out = {
        "metric": "llama2_7b_q4_0_prefill512_decode256_tok_s", "value": round((S + ND) / (pre + dec) * 1e3, 1), "unit": "tok/s", "n_gpus": world,
        "steps": args.steps, "warmup": max(args.warmup, 1), "ms_per_step": round(pre + dec, 3), "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
        "dtype": "int8 block dot (Q4_0 x Q8_0) + fp32 accumulate, fp16 KV", "data": "synthetic",
        "config": {"workload": "LLaMA-2-7B Q4_0 (BASELINE configs[1]): one request, prefill 512 + greedy decode 256, random-init .mllm; every Linear sharded over the ranks",
                   "l2": "weights 3.7 GB / N per decode step >> 126 MB L2 (inputs larger than L2)", "parallelism": f"tp{world}", "exchange": exchange},
        "decode_step": {"ms": round(step_ms, 4), "GB_per_rank": round(per_rank / 1e9, 4), "GBs_per_rank": round(per_rank / (step_ms * 1e-3) / 1e9, 1),
                        "frac_of_hbm_peak": round(per_rank / (step_ms * 1e-3) / 1e9 / peak, 4), "exchanges_per_token": 0 if world == 1 else 2 * cfg["layers"],
                        "exchange_bytes": cfg["hidden"] * 4},
        "prefill": {"ms": round(pre, 3), "tok_s": round(S / pre * 1e3, 1)},
        "e2e": {"value": round((S + ND) / wall * 1e3, 1), "unit": "tok/s", "h2d_bytes_per_step": S * 4, "d2h_bytes_per_step": ND * 4 + 8, "ms_per_step": round(wall, 2)},
        "gpu_launches": int(launches), "clocks": clocks, "tokens_head": toks[:8],
    }
    print(json.dumps(out), flush=True)


def reference_arm(args, rank, world):
    if rank != 0:
        return
    cfg = mf.config(MODEL, cache_limit=LIMIT)
    path = os.environ.get("MLLM_BENCH_MODEL", f"/tmp/mllm_bench_{MODEL}_q4_0.mllm")
    if not os.path.exists(path):
        import torch
        make_model_file(cfg, path, device="cuda" if torch.cuda.is_available() else "cpu")
    cp = args.cpu_prefill or (512 if args.steps <= 2 else 128 if args.steps <= 6 else 64)
    cd = args.cpu_decode or (32 if args.steps <= 2 else 16 if args.steps <= 6 else 8)
    r = cpu_arm(cfg, path, cp, cd, reps=args.steps)
    ms = r["seconds"] * 1e3
    out = {"impl": "reference", "metric": METRIC, "value": round(r["value"], 3), "unit": "tok/s", "n_gpus": world, "steps": args.steps,
           "warmup": args.warmup, "ms_per_step": round(ms, 1), "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
           "dtype": "int8 block dot (Q4_0 x Q8_0) + fp32 accumulate, fp16 KV", "data": "synthetic",
           "config": {"workload": "LLaMA-2-7B Q4_0 (BASELINE configs[1]) on the host CPU: " + r["sample"]},
           "cpu_baseline": {k: r[k] for k in ("value", "unit", "cores", "host_cpus", "kind", "sample", "prefill_tok_s", "decode_tok_s")},
           "prefill_tok_s": round(r["prefill_tok_s"], 3), "decode_tok_s": round(r["decode_tok_s"], 3) if r["decode_tok_s"] else None,
           "e2e": {"value": round(r["value"], 3), "unit": "tok/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(out), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--config", default=os.environ.get("MLLM_BENCH_CONFIG", "c2"), choices=["c2", "c5", "tp"],
                    help="c2: LLaMA-2-7B single request (the headline metric, default); c5: Qwen1.5-1.8B batched decode, data-parallel; tp: the 7B request with tensor-parallel Linears over the ranks")
    ap.add_argument("--batch", type=int, default=int(os.environ.get("MLLM_BENCH_BATCH", "64")), help="c5: requests per GPU")
    ap.add_argument("--cpu-prefill", type=int, default=0, help="CPU sample: prompt tokens (0 = 512 when steps <= 2, else 128 / 64 so the run stays within minutes)")
    ap.add_argument("--cpu-decode", type=int, default=0, help="CPU sample: decode steps (0 = 32 / 16 / 8 by the same rule)")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", 0))
    world = int(os.environ.get("WORLD_SIZE", 1))
    local_rank = int(os.environ.get("LOCAL_RANK", 0))
    if args.impl == "reference":
        return reference_arm(args, rank, world)
    if world > 1:
        import torch
        import torch.distributed as dist
        torch.cuda.set_device(local_rank)
        dist.init_process_group("nccl")
    if args.config == "tp":
        gpu_arm_tp(args, rank, world, local_rank)
    elif args.config == "c5":
        gpu_arm_c5(args, rank, world, local_rank)
    else:
        gpu_arm(args, rank, world, local_rank)
    if world > 1:
        import torch.distributed as dist
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
