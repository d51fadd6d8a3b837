"""Optional tensor-parallel LLaMA-family inference (north_star: "an optional column/row-parallel Linear for the 7B models uses NCCL
all-gather/all-reduce over NVLink").  One process per GPU (torch.distributed, backend nccl for the plumbing):

  * q / k / v and gate / up are COLUMN-parallel: rank r owns the weight rows of its heads / of its slice of the FFN width;
  * o_proj and down_proj are ROW-parallel: rank r owns the matching K-slice of every weight row (whole Q4_0 blocks; the FFN width is
    cut at multiples of 128 columns, unevenly when ranks do not divide it: 11008 = 86 x 128 -> 11,11,...,10 units at 8 ranks), and the
    partial sums are all-reduced before the residual add;
  * attention is local: each rank keeps the KV cache of its own kv heads; embedding, norms and lm_head are replicated (the lm_head
    stream is 17 us of the 7B step -- an all-gather of logits would cost more than it saves at this size), so every rank picks the
    same token without a further exchange.

The all-reduce is where the time goes: a decode step has 2 per layer, each 16 KiB for the 7B.
  prefill (hundreds of rows, MBs): torch.distributed.all_reduce -> NCCL over NVLink / NVSwitch, bandwidth-bound;
  decode (one row): one kernel per rank over CUDA-IPC peer memory (csrc/p2p.cu: flag exchange + direct peer loads + the residual add
  fused in), launched with programmatic dependent launch so it sits resident behind the producing GEMV and the next GEMV starts
  streaming its weights during the exchange; MLLM_TP_NCCL=1 keeps NCCL there for comparison (B200 x 2, LLaMA-2-7B: 1.44 ms/token
  against 1.97 with NCCL and 1.45 on one GPU).
The whole decode step (embedding, 7 launches per layer, head, pick) is captured in a CUDA graph per rank and replayed in lockstep.

The reference has no counterpart (one process, CPUBackend threads); the arithmetic per rank is the reference's Linear / RMSNorm / RoPE
/ attention on slices, so the only numerical difference to the 1-GPU path is the order of the fp32 partial sums of o / down.
"""
import ctypes as C
import os

import numpy as np

from . import capi, mllm_file as mf


# ------------------------------------------------------------------------------------------------------------------ sharding
def split_units(total, unit, world):
    """[(lo, hi)] per rank: `total` cut at multiples of `unit`, sizes differing by at most one unit."""
    if total % unit:
        raise ValueError(f"{total} is not a multiple of {unit}")
    n = total // unit
    if n < world:
        raise ValueError(f"cannot give {world} ranks at least {unit} of {total}")
    base, extra = divmod(n, world)
    out, lo = [], 0
    for r in range(world):
        hi = lo + (base + (1 if r < extra else 0)) * unit
        out.append((lo, hi))
        lo = hi
    return out


def shard_plan(cfg, world):
    """Per rank: heads [q0,q1), kv heads [k0,k1), FFN columns [f0,f1)."""
    Hq, Hkv = cfg["heads"], cfg["kv_heads"]
    if Hkv % world or Hq % world:
        raise ValueError(f"heads {Hq} / kv heads {Hkv} do not divide over {world} ranks")
    ffn = split_units(cfg["ffn"], 128, world)  # the decode GEMV wants K % 128 == 0
    return [dict(q=(r * Hq // world, (r + 1) * Hq // world), kv=(r * Hkv // world, (r + 1) * Hkv // world), ffn=ffn[r]) for r in range(world)]


def rows(w, lo, hi):
    """Column-parallel: rows [lo, hi) of a block-quantised [N][K/32*18] (or fp32 [N]) parameter."""
    return np.ascontiguousarray(w[lo:hi])


def k_slice_q4_0(w, N, K, lo, hi):
    """Row-parallel: columns [lo, hi) (multiples of 32) of every row of a Q4_0 [N][K/32][18] parameter."""
    if lo % 32 or hi % 32:
        raise ValueError("row-parallel cuts must fall on Q4_0 block boundaries")
    return np.ascontiguousarray(np.asarray(w).reshape(N, K // 32, 18)[:, lo // 32:hi // 32, :]).reshape(N, (hi - lo) // 32 * 18)


def shard_tensors(tensors, cfg, rank, world):
    """name -> numpy array of this rank's part of every parameter (replicated ones whole)."""
    n = mf.Names(cfg["rope_type"])
    D, H, F = cfg["head_dim"], cfg["hidden"], cfg["ffn"]
    p = shard_plan(cfg, world)[rank]
    (q0, q1), (k0, k1), (f0, f1) = p["q"], p["kv"], p["ffn"]
    V = cfg["vocab"]
    out = {n.embed: np.asarray(tensors[n.embed]).reshape(V, -1), n.norm: tensors[n.norm], n.lm_head: np.asarray(tensors[n.lm_head]).reshape(V, -1)}
    for L in range(cfg["layers"]):
        out[n.layer(L, "attn_norm")] = tensors[n.layer(L, "attn_norm")]
        out[n.layer(L, "ffn_norm")] = tensors[n.layer(L, "ffn_norm")]
        for role, (a, b) in (("wq", (q0, q1)), ("wk", (k0, k1)), ("wv", (k0, k1))):
            out[n.layer(L, role)] = rows(np.asarray(tensors[n.layer(L, role)]).reshape(-1, H // 32 * 18), a * D, b * D)
            if cfg["qkv_bias"]:
                out[n.layer(L, "b" + role[1])] = rows(np.asarray(tensors[n.layer(L, "b" + role[1])]).reshape(-1), a * D, b * D)
        out[n.layer(L, "wo")] = k_slice_q4_0(tensors[n.layer(L, "wo")], H, cfg["heads"] * D, q0 * D, q1 * D)
        out[n.layer(L, "w_gate")] = rows(np.asarray(tensors[n.layer(L, "w_gate")]).reshape(F, H // 32 * 18), f0, f1)
        out[n.layer(L, "w_up")] = rows(np.asarray(tensors[n.layer(L, "w_up")]).reshape(F, H // 32 * 18), f0, f1)
        out[n.layer(L, "w_down")] = k_slice_q4_0(tensors[n.layer(L, "w_down")], H, F, f0, f1)
    return out


def rope_tables(rope_type, limit, D, theta):
    """sin / cos [limit][D] as CPURoPE builds them (op/CPURoPE.cpp:75-128): theta_i = 1 / base^(2i/D) in double -> float, angle = s *
    theta_i in float, libm sinf / cosf; LLaMA (type 2) interleaves pairs, HF (type 4) splits halves."""
    libm = C.CDLL("libm.so.6")
    libm.sinf.restype = libm.cosf.restype = C.c_float
    libm.sinf.argtypes = libm.cosf.argtypes = [C.c_float]
    half = D // 2
    th = np.array([np.float32(1.0 / (float(theta) ** (2.0 * i / D))) for i in range(half)], dtype=np.float32)
    s = np.zeros((limit, D), dtype=np.float32)
    c = np.zeros((limit, D), dtype=np.float32)
    for pos in range(limit):
        ang = (np.float32(pos) * th).astype(np.float32)
        sn = np.array([libm.sinf(float(a)) for a in ang], dtype=np.float32)
        cs = np.array([libm.cosf(float(a)) for a in ang], dtype=np.float32)
        if rope_type == 2:  # LLAMAROPE: (2i, 2i+1) share angle i
            s[pos, 0::2], s[pos, 1::2], c[pos, 0::2], c[pos, 1::2] = sn, sn, cs, cs
        else:  # HFHUBROPE: i and i + D/2 share angle i
            s[pos, :half], s[pos, half:], c[pos, :half], c[pos, half:] = sn, sn, cs, cs
    return s, c


# ------------------------------------------------------------------------------------------------------------------ peer memory
class _Raw:
    """A device range that is not a torch tensor (cudaMalloc'ed, IPC-shared) with the few attributes capi.gemv_fused_ex reads."""

    def __init__(self, ptr, shape, esz=4):
        self.ptr, self.shape, self.esz = ptr, tuple(shape), esz

    def data_ptr(self):
        return self.ptr

    def element_size(self):
        return self.esz


class PeerRegion:
    """One cudaMalloc'ed region per rank, mapped into every peer through CUDA IPC.  Layout: n_sites x [world flag words, padded to
    256 B] then n_sites x [n floats]."""

    def __init__(self, dist, rank, world, n_sites, n):
        import torch
        self.rank, self.world, self.n_sites, self.n = rank, world, n_sites, n
        self.flag_bytes = 256
        self.buf_bytes = (n * 4 + 255) // 256 * 256
        self.bytes = n_sites * (self.flag_bytes + self.buf_bytes)
        lib = capi.lib()
        p = C.c_void_p()
        capi.check(lib.mllmcu_malloc(C.byref(p), self.bytes), "malloc")
        self.own = p.value
        capi.check(lib.mllmcu_memset(self.own, 0, self.bytes, None), "memset")
        torch.cuda.synchronize()
        h = (C.c_ubyte * 64)()
        capi.check(lib.mllmcu_ipc_get_handle(self.own, h), "ipc_get_handle")
        handles = [None] * world
        dist.all_gather_object(handles, bytes(h))
        self.bases = (C.c_void_p * world)()
        self.opened = []
        for r in range(world):
            if r == rank:
                self.bases[r] = self.own
                continue
            q = C.c_void_p()
            hb = (C.c_ubyte * 64).from_buffer_copy(handles[r])
            capi.check(lib.mllmcu_ipc_open(hb, C.byref(q)), "ipc_open")
            self.bases[r] = q.value
            self.opened.append(q.value)
        self.epoch = torch.zeros(4, dtype=torch.int32, device="cuda")
        self.ticket = torch.zeros(4, dtype=torch.int32, device="cuda")  # the fused GEMV tail's "last CTA" counter (self-resetting)
        dist.barrier()

    def flag_offset(self, site):
        return site * self.flag_bytes

    def buf_offset(self, site):
        return self.n_sites * self.flag_bytes + site * self.buf_bytes

    def buf(self, site, rows, cols):
        return _Raw(self.own + self.buf_offset(site), (rows, cols))

    def bump(self):
        capi.check(capi.lib().mllmcu_p2p_bump(self.epoch.data_ptr(), capi._stream()), "p2p_bump")

    def allreduce(self, site, residual, out, n, pdl=False):
        capi.check(capi.lib().mllmcu_p2p_allreduce(self.world, self.rank, self.bases, self.buf_offset(site), self.flag_offset(site), self.epoch.data_ptr(),
                                                   capi._ptr(residual), capi._ptr(out), n, int(pdl), capi._stream()), "p2p_allreduce")

    def close(self):
        lib = capi.lib()
        for q in self.opened:
            lib.mllmcu_ipc_close(q)
        lib.mllmcu_free(self.own)
        self.opened, self.own = [], None


# ------------------------------------------------------------------------------------------------------------------ the model
class TPLlama:
    """LLaMA-family decoder (models/llama/modeling_llama.hpp:29-115 op for op) with its Linears sharded over `world` ranks.
    world == 1 runs the same code without any exchange (the baseline the sharded logits are compared with)."""

    def __init__(self, cfg, path, rank=0, world=1, dist=None, batch=1, p2p=None):
        import torch
        self.torch, self.dist = torch, dist
        self.cfg, self.rank, self.world, self.B = cfg, rank, world, batch
        if cfg.get("wtype", mf.Q4_0) != mf.Q4_0:
            raise ValueError("tensor parallelism is built for Q4_0 models (Q4_K super-blocks of 256 do not divide 11008 / ranks)")
        self.dev = torch.device("cuda", torch.cuda.current_device())
        self.plan = shard_plan(cfg, world)[rank]
        tensors, _ = mf.read_mllm(path)
        sh = shard_tensors(tensors, cfg, rank, world)
        self.w = {k: torch.from_numpy(np.array(v, copy=True)).to(self.dev) for k, v in sh.items()}
        n = mf.Names(cfg["rope_type"])
        self.n = n
        D, H = cfg["head_dim"], cfg["hidden"]
        self.D, self.H = D, H
        self.Hq, self.Hkv = self.plan["q"][1] - self.plan["q"][0], self.plan["kv"][1] - self.plan["kv"][0]
        self.F = self.plan["ffn"][1] - self.plan["ffn"][0]
        self.limit = cfg["cache_limit"]
        s, c = rope_tables(cfg["rope_type"], self.limit, D, cfg["rope_theta"])
        self.sin, self.cos = torch.from_numpy(s).to(self.dev), torch.from_numpy(c).to(self.dev)
        self.kc = [torch.zeros((batch, self.limit, self.Hkv, D), dtype=torch.float16, device=self.dev) for _ in range(cfg["layers"])]
        self.vc = [torch.zeros((batch, self.limit, self.Hkv, D), dtype=torch.float16, device=self.dev) for _ in range(cfg["layers"])]
        self.ctl = torch.zeros(4, dtype=torch.int32, device=self.dev)
        self.pos = 0
        self.use_p2p = (world > 1 and os.environ.get("MLLM_TP_NCCL", "0") != "1") if p2p is None else (p2p and world > 1)
        self.peer = PeerRegion(dist, rank, world, 2 * cfg["layers"], batch * H) if self.use_p2p else None
        self.graph = None
        self.launches_per_step = 0
        self.pdl = os.environ.get("MLLM_CUDA_NO_PDL", "0") in ("", "0")
        # MLLM_TP_FUSED=1: the exchange as the producing GEMV's tail (one launch instead of two).  Measured slower on B200 at TP 2 (1.69
        # against 1.44 ms/token): the next GEMV, released at the producer's START, is already streaming its weights when the tail
        # exchanges, and the flag round trip queues behind that traffic; as a separate kernel the exchange overlaps the next GEMV's
        # launch ramp instead.  Kept selectable for that comparison; default off.
        self.fused_exchange = os.environ.get("MLLM_TP_FUSED", "0") == "1"

    # ---- parameters of layer L
    def _lw(self, L, role):
        return self.w[self.n.layer(L, role)]

    def _bias(self, L, role):
        return self.w[self.n.layer(L, role)] if self.cfg["qkv_bias"] else None

    def _reduce(self, t):
        if self.world > 1:
            self.dist.all_reduce(t)
        return t

    # ---- prefill: op by op, NCCL all-reduce of the row-parallel partial sums
    def prefill(self, ids):
        """ids: float/int array [B, S]; fills the caches from position self.pos; returns the last-position logits [B, V] (torch)."""
        torch, cfg = self.torch, self.cfg
        ids = np.asarray(ids, dtype=np.float32).reshape(self.B, -1)
        B, S = ids.shape
        D, H, Hq, Hkv = self.D, self.H, self.Hq, self.Hkv
        eps, rt = cfg["rms_eps"], cfg["rope_type"]
        p0 = self.pos
        if p0 + S > self.limit:
            raise ValueError("KV cache limit")
        x = capi.embedding_q4_0(self.w[self.n.embed], torch.from_numpy(ids.reshape(-1)).to(self.dev), H)  # [B*S, H]
        lib = capi.lib()
        for L in range(cfg["layers"]):
            h = capi.rmsnorm(x, self._lw(L, "attn_norm"), eps)
            q = capi.linear(self._lw(L, "wq"), h, self._bias(L, "bq"))
            k = capi.linear(self._lw(L, "wk"), h, self._bias(L, "bk"))
            v = capi.linear(self._lw(L, "wv"), h, self._bias(L, "bv"))
            q = capi.rope(rt, q.view(B, S, Hq, D), p0, self.sin, self.cos)
            kc, vc = self.kc[L], self.vc[L]
            capi.rope(rt, k.view(B, S, Hkv, D), p0, self.sin, self.cos, out_f16=True, out=kc[:, p0:], out_bstride=kc.stride(0), out_sstride=kc.stride(1))
            for b in range(B):  # v rows -> fp16 cache rows (the KVCache append, op/CPUKVCache.cpp:137-247)
                capi.check(lib.mllmcu_f32_to_f16_rows(v[b * S:(b + 1) * S].data_ptr(), vc[b, p0:].data_ptr(), S, Hkv * D, Hkv * D, Hkv * D, capi._stream()), "f32_to_f16")
            a = capi.flash_attn_f16kv(q, kc[:, :p0 + S], vc[:, :p0 + S], causal=True)
            o = self._reduce(capi.linear(self._lw(L, "wo"), a.view(B * S, Hq * D)))
            x = capi.add(x, o)
            h = capi.rmsnorm(x, self._lw(L, "ffn_norm"), eps)
            g = capi.linear(self._lw(L, "w_gate"), h)
            u = capi.linear(self._lw(L, "w_up"), h)
            d = self._reduce(capi.linear(self._lw(L, "w_down"), capi.silu_mul(g, u)))
            x = capi.add(x, d)
        last = x.view(B, S, H)[:, S - 1].contiguous()
        logits = capi.linear(self.w[self.n.lm_head], capi.rmsnorm(last, self.w[self.n.norm], eps))
        self.pos = p0 + S
        return logits

    # ---- decode: the fused single-token kernels, 7 launches per layer, peer-memory all-reduce with the residual add fused in
    def _step_body(self, ids_dev):
        torch, cfg = self.torch, self.cfg
        D, H, Hq, B = self.D, self.H, self.Hq, self.B
        eps, rt = cfg["rms_eps"], cfg["rope_type"]
        n_launch = 0
        if self.peer:
            self.peer.bump()
            n_launch += 1
        x = capi.decode_embed(self.w[self.n.embed], ids_dev, H, self.ctl)  # ctl[0] += 1: the position of this token
        n_launch += 1
        for L in range(cfg["layers"]):
            bq = [self._bias(L, "bq"), self._bias(L, "bk"), self._bias(L, "bv")] if cfg["qkv_bias"] else None
            # programmatic dependent launch as in the 1-GPU plan (csrc/host/Module.cpp issue_step): layer 0's projection releases its
            # dependents only after its own wait (2), so everything behind it may read the position ahead of its own wait
            qkv = capi.gemv_fused([self._lw(L, "wq"), self._lw(L, "wk"), self._lw(L, "wv")], x, norm_w=self._lw(L, "attn_norm"), eps=eps, biases=bq,
                                  pdl=(2 if L == 0 else 1) if self.pdl else 0)
            a = capi.decode_attn_rope(rt, qkv, self.kc[L], self.vc[L], self.ctl, Hq, D, self.sin, self.cos, pdl=2 if self.pdl else 0)
            x = self._row_parallel(2 * L, [self._lw(L, "wo")], a, x, epilogue=0)
            x = self._row_parallel(2 * L + 1, [self._lw(L, "w_down")],
                                   capi.gemv_fused([self._lw(L, "w_gate"), self._lw(L, "w_up")], x, norm_w=self._lw(L, "ffn_norm"), eps=eps, epilogue=2, pdl=self.pdl), x, epilogue=0)
            n_launch += 4 + ((0 if (self.fused_exchange and self.B <= 2) else 2) if self.peer else (4 if self.world > 1 else 0))
        logits = capi.gemv_fused([self.w[self.n.lm_head]], x, norm_w=self.w[self.n.norm], eps=eps, pdl=self.pdl)
        nxt = capi.decode_argmax(logits)
        self.launches_per_step = n_launch + 2
        return logits, nxt

    def _row_parallel(self, site, Ws, a, residual, epilogue):
        """x + all_reduce(W_slice . a): 1 rank -> the GEMV's own residual epilogue; P2P -> partial into the shared site, one exchange
        kernel with the residual add; NCCL -> partial, all_reduce, add."""
        torch = self.torch
        if self.world == 1:
            return capi.gemv_fused(Ws, a, epilogue=1, residual=residual, pdl=self.pdl)
        if self.peer:
            y = self.peer.buf(site, self.B, self.H)
            out = torch.empty_like(residual)
            if self.fused_exchange and self.B <= 2:  # Linear + exchange + residual in ONE launch: the GEMV's last CTA does the exchange
                capi.gemv_fused_ex(Ws, a, y, pdl=self.pdl, xchg=dict(world=self.world, rank=self.rank, bases=self.peer.bases, buf_offset=self.peer.buf_offset(site),
                                                                       flag_offset=self.peer.flag_offset(site), epoch=self.peer.epoch, ticket=self.peer.ticket,
                                                                       residual=residual, out=out))
                return out
            capi.gemv_fused_ex(Ws, a, y, pdl=self.pdl)
            self.peer.allreduce(site, residual, out, self.B * self.H, pdl=self.pdl)
            return out
        part = capi.gemv_fused(Ws, a, pdl=self.pdl)
        self.dist.all_reduce(part)
        return capi.add(residual, part)

    def decode(self, first_ids, n_steps, graph=True):
        """Greedy steps from `first_ids` [B] (the token picked after the prefill).  Returns (tokens [n_steps][B], last logits)."""
        torch = self.torch
        self.ctl[0] = self.pos - 1
        ids = torch.as_tensor(np.asarray(first_ids, dtype=np.float32).reshape(self.B), device=self.dev).clone()
        toks = []
        if not graph:
            for _ in range(n_steps):
                logits, nxt = self._step_body(ids)
                ids = nxt
                toks.append(nxt.cpu().numpy().astype(np.int64))
                self.pos += 1
            return np.stack(toks), logits
        if self.graph is None:
            # one eager step first: kernels allocate their scratch lazily, never while a stream is capturing
            self._ids = ids
            logits, nxt = self._step_body(self._ids)
            self._ids.copy_(nxt)
            toks.append(nxt.cpu().numpy().astype(np.int64))
            self.pos += 1
            n_steps -= 1
            torch.cuda.synchronize()
            if self.dist is not None and self.world > 1:
                self.dist.barrier()
            self.graph = torch.cuda.CUDAGraph()
            self._tok_log = torch.zeros((self.limit, self.B), dtype=torch.float32, device=self.dev)
            with torch.cuda.graph(self.graph):
                self._logits, nxt = self._step_body(self._ids)
                self._ids.copy_(nxt)
                self._tok_log.index_copy_(0, self.ctl[:1].to(torch.int64), nxt.view(1, self.B))  # token log row = position
            # the capture did not run anything: ctl / epoch are where the eager step left them
        else:
            self._ids.copy_(ids)
        p_first = self.pos
        for _ in range(n_steps):
            self.graph.replay()
            self.pos += 1
        torch.cuda.synchronize()
        if n_steps > 0:
            got = self._tok_log[p_first:p_first + n_steps].cpu().numpy().astype(np.int64)
            toks.extend(list(got))
            logits = self._logits
        return np.stack(toks), logits

    def clear(self):
        self.pos = 0

    def close(self):
        self.graph = None
        if self.peer:
            self.peer.close()
            self.peer = None
