"""ctypes access to the CPU checkers.  TEST INFRASTRUCTURE ONLY.

`port()` -> oracle/liboracle.so (plain-C restatement, oracle.c)
`ref()`  -> oracle/_ref/libmllm_ref.so (the reference's own sources, compiled by
            oracle/Makefile) or None when it has not been built.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl
reference legs may import this module; mllm_b200/ never does.
"""
import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
_port = None
_ref = None

Q4_0, Q8_0 = 2, 8  # mllm/Types.hpp:64-68 DataType values
Q4_0_BYTES, Q8_0_BYTES = 18, 34
Q4_K, Q6_K = 12, 14
Q4_K_BYTES, Q6_K_BYTES, Q8_K_BYTES = 144, 210, 292


def build():
    """(Re)build liboracle.so and, where /root/reference exists, _ref/libmllm_ref.so."""
    subprocess.run(["make", "-s", "-C", HERE], check=True)


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


class Lib:
    """Uniform numpy front-end over either library (prefix 'orc_' or 'ref_')."""

    def __init__(self, path, prefix):
        self.lib = C.CDLL(path)
        self.prefix = prefix
        self.path = path
        L = self.lib
        for name in ("fp32_to_fp16",):
            f = getattr(L, prefix + name)
            f.restype = C.c_uint16
            f.argtypes = [C.c_float]
        f = getattr(L, prefix + "fp16_to_fp32")
        f.restype = C.c_float
        f.argtypes = [C.c_uint16]
        L.mr_create.restype = C.c_void_p
        L.mr_create.argtypes = [C.c_void_p, C.c_void_p]
        L.mr_destroy.argtypes = [C.c_void_p]
        L.mr_set_layer.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p]
        L.mr_set_global.argtypes = [C.c_void_p, C.c_int, C.c_void_p]
        L.mr_reset.argtypes = [C.c_void_p]
        L.mr_cache_len.argtypes = [C.c_void_p]
        L.mr_forward.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p]
        L.mr_forward.restype = C.c_int

    def fn(self, name):
        return getattr(self.lib, self.prefix + name)

    # ---- rows
    def quantize_q8_0(self, x):
        x = np.ascontiguousarray(x, dtype=np.float32)
        rows, k = x.reshape(-1, x.shape[-1]).shape
        out = np.empty((rows, k // 32 * Q8_0_BYTES), dtype=np.uint8)
        f = self.fn("quantize_row_q8_0")
        xr = x.reshape(rows, k)
        for r in range(rows):
            f(_p(xr[r]), _p(out[r]), C.c_int(k))
        return out

    def quantize_q4_0(self, x):
        x = np.ascontiguousarray(x, dtype=np.float32)
        rows, k = x.reshape(-1, x.shape[-1]).shape
        out = np.empty((rows, k // 32 * Q4_0_BYTES), dtype=np.uint8)
        f = self.fn("quantize_row_q4_0")
        xr = x.reshape(rows, k)
        for r in range(rows):
            f(_p(xr[r]), _p(out[r]), C.c_int(k))
        return out

    def dequantize_q4_0(self, w, k):
        w = np.ascontiguousarray(w, dtype=np.uint8).reshape(-1, k // 32 * Q4_0_BYTES)
        out = np.empty((w.shape[0], k), dtype=np.float32)
        f = self.fn("dequantize_row_q4_0")
        for r in range(w.shape[0]):
            f(_p(w[r]), _p(out[r]), C.c_int(k))
        return out

    def vec_dot_q4_0_q8_0(self, wrow, xrow, k):
        s = C.c_float(0)
        self.fn("vec_dot_q4_0_q8_0")(C.c_int(k), C.byref(s), _p(wrow), _p(xrow))
        return np.float32(s.value)

    def linear(self, W, x, bias=None, wtype=Q4_0, out_fp16=False):
        """x [M,K] fp32, W [N, K/32*blk] uint8 -> y [M,N] fp32 (or uint16 holding fp16)."""
        x = np.ascontiguousarray(x, dtype=np.float32)
        M, K = x.shape
        N = W.shape[0]
        y = np.empty((M, N), dtype=np.uint16 if out_fp16 else np.float32)
        b = _p(np.ascontiguousarray(bias, dtype=np.float32)) if bias is not None else None
        rc = self.fn("linear")(C.c_int(wtype), _p(W), _p(x), b, _p(y), M, N, K, K, N, int(out_fp16))
        assert rc == 0
        return y

    def silu(self, x):
        x = np.ascontiguousarray(x, dtype=np.float32)
        y = np.empty_like(x)
        if self.prefix == "orc_":
            self.fn("silu")(_p(x), _p(y), C.c_size_t(x.size), 8)
        else:
            self.fn("silu")(_p(x), _p(y), C.c_size_t(x.size))
        return y

    def flash_attn_f16kv(self, q, k, v, causal=True, threads=1):
        """q [Sq,Hq,D] fp32; k,v [Sk,Hkv,D] uint16 (fp16 bits) contiguous -> o [Sq,Hq,D]."""
        q = np.ascontiguousarray(q, dtype=np.float32)
        k = np.ascontiguousarray(k, dtype=np.uint16)
        v = np.ascontiguousarray(v, dtype=np.uint16)
        Sq, Hq, D = q.shape
        Sk, Hkv, _ = k.shape
        o = np.empty_like(q)
        if self.prefix == "orc_":
            self.fn("flash_attn_f16kv")(_p(q), _p(k), _p(v), _p(o), Sq, Sk, Hq, Hkv, D, int(causal), C.c_size_t(Hkv * D))
        else:
            self.fn("flash_attn_f16kv")(_p(q), _p(k), _p(v), _p(o), Sq, Sk, Hq, Hkv, D, int(causal), threads)
        return o

    # ---- rows a4/a5/a7/a15/a17: available with both prefixes (restatement and compiled reference)
    def quantize_q8_K(self, x):
        x = np.ascontiguousarray(x, dtype=np.float32)
        rows, k = x.reshape(-1, x.shape[-1]).shape
        out = np.zeros((rows, k // 256 * Q8_K_BYTES), dtype=np.uint8)
        f = self.fn("quantize_row_q8_K")
        xr = x.reshape(rows, k)
        for r in range(rows):
            f(_p(xr[r]), _p(out[r]), C.c_int(k))
        return out

    def dequantize_kq(self, w, k, wtype=Q4_K):
        bs = Q4_K_BYTES if wtype == Q4_K else Q6_K_BYTES
        w = np.ascontiguousarray(w, dtype=np.uint8).reshape(-1, k // 256 * bs)
        out = np.empty((w.shape[0], k), dtype=np.float32)
        f = self.fn("dequantize_row_q4_K" if wtype == Q4_K else "dequantize_row_q6_K")
        for r in range(w.shape[0]):
            f(_p(w[r]), _p(out[r]), C.c_int(k))
        return out

    def vec_dot_kq(self, wrow, xrow, k, wtype=Q4_K):
        s = C.c_float(0)
        if wtype == Q4_K:
            self.fn("vec_dot_q4_K_q8_K")(C.c_int(k), C.byref(s), _p(wrow), _p(xrow))
        elif self.prefix == "orc_":
            self.fn("vec_dot_q6_K_q8_K")(C.c_int(k), C.byref(s), _p(wrow), _p(xrow), None)
        else:
            self.fn("vec_dot_q6_K_q8_K")(C.c_int(k), C.byref(s), _p(wrow), _p(xrow))
        return np.float32(s.value)

    def linear_kq(self, W, x, bias=None, wtype=Q4_K):
        x = np.ascontiguousarray(x, dtype=np.float32)
        M, K = x.shape
        N = W.shape[0]
        y = np.empty((M, N), dtype=np.float32)
        b = _p(np.ascontiguousarray(bias, dtype=np.float32)) if bias is not None else None
        rc = self.fn("linear_kq")(C.c_int(wtype), _p(W), _p(x), b, _p(y), M, N, K, K, N)
        assert rc == 0
        return y

    def softmax(self, x, causal=False, old_dim=0):
        """x [outer, S, n] -> CPUSoftMax over the last dim"""
        x = np.ascontiguousarray(x, dtype=np.float32)
        outer, S, n = x.shape
        y = np.empty_like(x)
        self.fn("softmax")(_p(x), _p(y), outer, S, n, int(causal), old_dim)
        return y

    def gelu_table(self, kind):
        t = np.empty(65536, dtype=np.uint16)
        if self.prefix == "orc_":
            self.fn("gelu_table")(kind, _p(t))
        else:
            self.fn("gelu_table")(kind, _p(t))
        return t

    def conv2d_valid(self, x, w, bias, stride):
        x = np.ascontiguousarray(x, dtype=np.float32)
        w = np.ascontiguousarray(w, dtype=np.float32)
        B, Cc, H, W = x.shape
        O, _, kh, kw = w.shape
        sh, sw = stride
        y = np.empty((B, O, (H - kh) // sh + 1, (W - kw) // sw + 1), dtype=np.float32)
        b = _p(np.ascontiguousarray(bias, dtype=np.float32)) if bias is not None else None
        self.fn("conv2d_valid")(_p(x), _p(w), b, _p(y), B, Cc, H, W, O, kh, kw, sh, sw)
        return y


class PortOnly:
    """Functions that exist only in the restatement (no compilable reference twin)."""

    def __init__(self, lib):
        self.L = lib.lib

    def rmsnorm(self, x, w, eps, add_unit_offset=False):
        x = np.ascontiguousarray(x, dtype=np.float32)
        rows, dim = x.reshape(-1, x.shape[-1]).shape
        y = np.empty_like(x)
        w = np.ascontiguousarray(w, dtype=np.float32)
        self.L.orc_rmsnorm(_p(x), _p(w), _p(y), rows, dim, C.c_float(eps), int(add_unit_offset))
        return y

    def layernorm(self, x, w, b, eps):
        x = np.ascontiguousarray(x, dtype=np.float32)
        rows, dim = x.reshape(-1, x.shape[-1]).shape
        y = np.empty_like(x)
        self.L.orc_layernorm(_p(x), _p(w) if w is not None else None, _p(b) if b is not None else None, _p(y), rows, dim, C.c_float(eps))
        return y

    def rope_tables(self, rtype, pos_max, dim, base):
        s = np.empty((pos_max, dim), dtype=np.float32)
        c = np.empty((pos_max, dim), dtype=np.float32)
        self.L.orc_rope_tables(rtype, pos_max, dim, C.c_float(base), _p(s), _p(c))
        return s, c

    def rope_apply(self, rtype, x, pos0, sin_t, cos_t, rot=None, out_fp16=False):
        x = np.ascontiguousarray(x, dtype=np.float32)
        S, H, D = x.shape
        rot = D if rot is None else rot
        out = np.empty((S, H, D), dtype=np.uint16 if out_fp16 else np.float32)
        self.L.orc_rope_apply(rtype, _p(x), _p(out), S, H, D, rot, pos0, _p(sin_t), _p(cos_t), sin_t.shape[1], int(out_fp16), C.c_size_t(H * D))
        return out

    def embedding_q4_0(self, table, ids, hidden):
        ids = np.ascontiguousarray(ids, dtype=np.float32).reshape(-1)
        out = np.zeros((ids.size, hidden), dtype=np.float32)
        self.L.orc_embedding_q4_0(_p(table), _p(ids), _p(out), ids.size, hidden)
        return out

    def block_isums_q4_0_q8_0(self, wrow, xrow, k):
        out = np.empty(k // 32, dtype=np.int32)
        self.L.orc_block_isums_q4_0_q8_0(k, _p(wrow), _p(xrow), _p(out))
        return out

    def gelu(self, table, x):
        x = np.ascontiguousarray(x, dtype=np.float32)
        y = np.empty_like(x)
        self.L.orc_gelu_lookup(_p(table), _p(x), _p(y), C.c_size_t(x.size))
        return y

    def mrope_tables(self, pos, dim, base, sections):
        pos = np.ascontiguousarray(pos, dtype=np.float32)  # [3, S]
        S = pos.shape[1]
        sec = np.ascontiguousarray(sections, dtype=np.int32)
        s = np.empty((S, dim), dtype=np.float32)
        c = np.empty((S, dim), dtype=np.float32)
        self.L.orc_mrope_tables(_p(pos), S, dim, C.c_float(base), _p(sec), sec.size, _p(s), _p(c))
        return s, c

    def rope_vision(self, x, angles):
        x = np.ascontiguousarray(x, dtype=np.float32)
        angles = np.ascontiguousarray(angles, dtype=np.float32)
        S, H, D = x.shape
        out = np.empty_like(x)
        self.L.orc_rope_vision(_p(x), _p(angles), _p(out), S, H, D)
        return out

    def conv3d_valid(self, x, w, bias, stride, channels_last=False):
        x = np.ascontiguousarray(x, dtype=np.float32)
        w = np.ascontiguousarray(w, dtype=np.float32)
        B, Cc, T, H, W = x.shape
        O, _, kt, kh, kw = w.shape
        st, sh, sw = stride
        ot, oh, ow = (T - kt) // st + 1, (H - kh) // sh + 1, (W - kw) // sw + 1
        y = np.empty((B, ot, oh, ow, O) if channels_last else (B, O, ot, oh, ow), dtype=np.float32)
        b = _p(np.ascontiguousarray(bias, dtype=np.float32)) if bias is not None else None
        self.L.orc_conv3d_valid(_p(x), _p(w), b, _p(y), B, Cc, T, H, W, O, kt, kh, kw, st, sh, sw, int(channels_last))
        return y

    def q4_0_4x4_pack(self, rows, k):
        rows = np.ascontiguousarray(rows, dtype=np.uint8)
        out = np.empty_like(rows)
        self.L.orc_q4_0_4x4_pack(_p(rows), _p(out), rows.shape[0], k)
        return out

    def q4_0_4x4_unpack(self, packed, nrow, k):
        packed = np.ascontiguousarray(packed, dtype=np.uint8)
        out = np.empty((nrow, k // 32 * Q4_0_BYTES), dtype=np.uint8)
        self.L.orc_q4_0_4x4_unpack(_p(packed), _p(out), nrow, k)
        return out

    def block_isums_q4_K_q8_K(self, wrow, xrow, k):
        a = np.empty(k // 256, dtype=np.int32)
        b = np.empty(k // 256, dtype=np.int32)
        self.L.orc_block_isums_q4_K_q8_K(k, _p(wrow), _p(xrow), _p(a), _p(b))
        return a, b

    def block_lanes_q6_K_q8_K(self, wrow, xrow, k):
        lanes = np.empty((k // 256, 8), dtype=np.int32)
        s = C.c_float(0)
        self.L.orc_vec_dot_q6_K_q8_K(C.c_int(k), C.byref(s), _p(wrow), _p(xrow), _p(lanes))
        return lanes

    def fp32_to_fp16_bits(self, x):
        x = np.ascontiguousarray(x, dtype=np.float32)
        f = self.L.orc_fp32_to_fp16
        f.restype = C.c_uint16
        f.argtypes = [C.c_float]
        return np.array([f(float(v)) for v in x.reshape(-1)], dtype=np.uint16).reshape(x.shape)


def port():
    global _port
    if _port is None:
        so = os.path.join(HERE, "liboracle.so")
        if not os.path.exists(so):
            build()
        _port = Lib(so, "orc_")
        _port.only = PortOnly(_port)
    return _port


def ref():
    """The compiled reference, or None if oracle/_ref has not been built."""
    global _ref
    if _ref is None:
        so = os.path.join(HERE, "_ref", "libmllm_ref.so")
        if not os.path.exists(so):
            return None
        _ref = Lib(so, "ref_")
    return _ref


class ModelRunner:
    """CPU forward of a LLaMA-family model held in a dict name -> numpy array (see mllm_b200.mllm_file)."""

    ROLES = dict(attn_norm=0, wq=1, wk=2, wv=3, wo=4, bq=5, bk=6, bv=7, ffn_norm=8, w_gate=9, w_up=10, w_down=11)

    def __init__(self, lib, cfg, tensors, names, max_batch=1, cache_limit=None):
        self.lib = lib.lib
        self.cfg = cfg
        self.keep = tensors
        cache_limit = cache_limit or cfg["cache_limit"]
        ci = np.array([cfg["vocab"], cfg["hidden"], cfg["layers"], cfg["heads"], cfg["kv_heads"], cfg["head_dim"],
                       cfg["ffn"], cfg["rope_type"], cache_limit, max_batch, (Q4_K if cfg.get("wtype", Q4_0) == Q4_K else Q4_0),
                       cfg.get("pos_max", cache_limit)], dtype=np.int32)
        cf = np.array([cfg["rope_theta"], cfg["rms_eps"]], dtype=np.float32)
        self.h = C.c_void_p(self.lib.mr_create(_p(ci), _p(cf)))
        for L in range(cfg["layers"]):
            for role, idx in self.ROLES.items():
                nm = names.layer(L, role)
                if nm in tensors:
                    self.lib.mr_set_layer(self.h, L, idx, _p(tensors[nm]))
        self.lib.mr_set_global(self.h, 0, _p(tensors[names.embed]))
        self.lib.mr_set_global(self.h, 1, _p(tensors[names.norm]))
        self.lib.mr_set_global(self.h, 2, _p(tensors[names.lm_head]))

    def reset(self):
        self.lib.mr_reset(self.h)

    def set_port_attention(self, on):
        """Reference-kernel build only: swap attention for the restatement's (see model_runner.c mr_set_attention)."""
        self.lib.mr_set_attention(int(on))

    def forward(self, ids, dump=False):
        ids = np.ascontiguousarray(ids, dtype=np.float32)
        B, S = ids.shape
        logits = np.empty((B, self.cfg["vocab"]), dtype=np.float32)
        d = np.empty((self.cfg["layers"] + 1, B * S, self.cfg["hidden"]), dtype=np.float32) if dump else None
        rc = self.lib.mr_forward(self.h, _p(ids), B, S, _p(logits), _p(d) if dump else None)
        if rc != 0:
            raise RuntimeError("KV cache limit exceeded")
        return (logits, d) if dump else logits

    def __del__(self):
        try:
            self.lib.mr_destroy(self.h)
        except Exception:
            pass


# ---------------------------------------------------------------- the reference's whole CPU backend (oracle/cpuref.mk)
_cpuref = None
F32, F16 = 0, 1
BSHD, BHSD, BCTHW = 0, 1, 3


class _TDesc(C.Structure):
    _fields_ = [("data", C.c_void_p), ("shape", C.c_int * 5), ("ndim", C.c_int), ("dtype", C.c_int), ("ctype", C.c_int)]


def build_cpuref():
    """g++ over the reference's CPU-backend sources where they lie (no cmake) + ref_ops_shim.cpp; needs /root/reference."""
    subprocess.run(["make", "-s", "-j", str(os.cpu_count() or 4), "-f", os.path.join(HERE, "cpuref.mk")], check=True, cwd=HERE)


def cpuref():
    """oracle/_ref/libmllm_cpu_ref.so (the reference's Op classes and model graphs on MLLM_CPU), or None when not built."""
    global _cpuref
    if _cpuref is None:
        so = os.path.join(HERE, "_ref", "libmllm_cpu_ref.so")
        if not os.path.exists(so):
            return None
        L = C.CDLL(so)
        L.cpuref_set_threads.argtypes = [C.c_int]
        L.cpuref_op_create.restype = C.c_void_p
        L.cpuref_op_create.argtypes = [C.POINTER(C.c_char_p), C.POINTER(C.c_float), C.c_int, C.c_char_p, C.c_int]
        L.cpuref_op_load.argtypes = [C.c_void_p, C.c_char_p]
        L.cpuref_op_run.argtypes = [C.c_void_p, C.POINTER(_TDesc), C.c_int, C.c_int]
        L.cpuref_op_out_info.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(C.c_int)]
        L.cpuref_op_out_read.argtypes = [C.c_void_p, C.c_int, C.c_void_p]
        L.cpuref_op_clear_cache.argtypes = [C.c_void_p]
        L.cpuref_op_destroy.argtypes = [C.c_void_p]
        L.cpuref_model_create.restype = C.c_void_p
        L.cpuref_model_create.argtypes = [C.c_char_p, C.c_int, C.c_int, C.c_void_p, C.c_float, C.c_float, C.c_int]
        L.cpuref_model_load.argtypes = [C.c_void_p, C.c_char_p]
        L.cpuref_model_vocab.argtypes = [C.c_void_p]
        L.cpuref_model_forward.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_int]
        L.cpuref_model_clear.argtypes = [C.c_void_p]
        L.cpuref_model_generate.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p]
        L.cpuref_model_destroy.argtypes = [C.c_void_p]
        _cpuref = L
    return _cpuref


class RefOp:
    """One of the reference's own CPU Op classes (CPUBackend::opCreate), driven as test/cpu/CPUTest.hpp drives them.
    params: the OpParam map (must hold "type" = OpType, OpDefined.hpp:10-134); weights: dict suffix -> (array, dtype) written to
    a temporary .mllm and loaded through the reference's ParamLoader under the names `name + suffix`."""

    def __init__(self, params, name="op", weights=None, threads=4):
        self.L = cpuref()
        if self.L is None:
            raise RuntimeError("oracle/_ref/libmllm_cpu_ref.so not built")
        keys = (C.c_char_p * len(params))(*[k.encode() for k in params])
        vals = (C.c_float * len(params))(*[float(v) for v in params.values()])
        self.h = C.c_void_p(self.L.cpuref_op_create(keys, vals, len(params), name.encode(), threads))
        if not self.h:
            raise RuntimeError(f"reference CPUBackend has no creator for {params}")
        if weights is not None:
            import tempfile
            from mllm_b200 import mllm_file as mf  # the container writer only (file format), no device code
            with tempfile.TemporaryDirectory() as d:
                p = os.path.join(d, "w.mllm")
                mf.write_mllm(p, {name + s: np.ascontiguousarray(a) for s, (a, _) in weights.items()}, {name + s: dt for s, (_, dt) in weights.items()})
                self.L.cpuref_op_load(self.h, p.encode())
        else:
            self.L.cpuref_op_load(self.h, b"")

    def run(self, inputs, n_out=1):
        """inputs: list of (array, shape, dtype, ctype); shape = (batch, head, sequence, dimension) -- the array's MEMORY order is the
        ctype's (BSHD: [b][s][h][d]) -- or 5 numbers (batch, channel, time, height, width) for BCTHW.  Returns fp32 arrays shaped
        [batch, sequence, head, dimension] (or [b, c, t, h, w])."""
        descs = (_TDesc * len(inputs))()
        keep = []
        for i, (a, shape, dtype, ctype) in enumerate(inputs):
            a = np.ascontiguousarray(a)
            keep.append(a)
            descs[i].data = a.ctypes.data
            for j, s in enumerate(shape):
                descs[i].shape[j] = int(s)
            descs[i].ndim, descs[i].dtype, descs[i].ctype = len(shape), dtype, ctype
        rc = self.L.cpuref_op_run(self.h, descs, len(inputs), n_out)
        if rc != 0:
            raise RuntimeError(f"reference op failed at stage {rc}")
        outs = []
        for i in range(n_out):
            shp = (C.c_int * 5)()
            dt, ct = C.c_int(), C.c_int()
            nd = self.L.cpuref_op_out_info(self.h, i, shp, C.byref(dt), C.byref(ct))
            b, x1, x2, x3, x4 = list(shp)
            out = np.empty((b, x2, x1, x3) if nd == 4 else (b, x1, x2, x3, x4), dtype=np.float32)
            self.L.cpuref_op_out_read(self.h, i, _p(out))
            outs.append(out)
        return outs

    def clear_cache(self):
        self.L.cpuref_op_clear_cache(self.h)

    def close(self):
        if self.h:
            self.L.cpuref_op_destroy(self.h)
            self.h = None


class RefGraph:
    """The reference's own model graph (LLaMAModel / TinyLLaMAModel / QWenForCausalLM) on MLLM_CPU over a .mllm file.
    family: "llama2-7b", "tinyllama-1.1b", "qwen1.5-1.8b", or "llama-generic" with cfg dims (toy models)."""

    def __init__(self, cfg, path, family=None, threads=None):
        self.L = cpuref()
        if self.L is None:
            raise RuntimeError("oracle/_ref/libmllm_cpu_ref.so not built")
        if family is None:
            if cfg["name"] in ("llama2-7b", "tinyllama-1.1b", "qwen1.5-1.8b"):
                family = cfg["name"]
            else:  # toy shapes: LLaMAModel (no bias, eps 1e-6 hard-coded in LLaMABlock) or QWenForCausalLM (qkv bias, config eps)
                family = "qwen-generic" if (cfg["qkv_bias"] or abs(cfg["rms_eps"] - 1e-6) > 1e-12) else "llama-generic"
        dims = np.array([cfg["vocab"], cfg["hidden"], cfg["heads"], cfg["kv_heads"], cfg["ffn"], cfg["layers"], cfg.get("pos_max", cfg["cache_limit"])], dtype=np.int32)
        self.vocab = cfg["vocab"]
        # CPURoPE keeps its sin/cos tables in process-global statics and rebuilds them only when the rotated width grows or the pose
        # type changes (op/CPURoPE.cpp:13-18,207-224: one model per process).  A graph made after other RoPE users of this process
        # would silently rotate with THEIR tables: flip the pose type with a throw-away op so this model builds its own.
        other = 2 if cfg["rope_type"] == 4 else 4
        dummy = RefOp({"type": 8, "pose_type": other, "rope_theta": 10000.0, "max_position_embeddings": 4}, "dummy", None)
        dummy.run([(np.zeros((1, 1, 1, 8), dtype=np.float32), (1, 1, 1, 8), F32, BSHD)])
        dummy.close()
        self.h = C.c_void_p(self.L.cpuref_model_create(family.encode(), cfg["cache_limit"], cfg["rope_type"], _p(dims), cfg["rope_theta"], cfg["rms_eps"],
                                                       threads or (os.cpu_count() or 1)))
        if not self.h:
            raise RuntimeError(f"unknown model family {family}")
        self.L.cpuref_model_load(self.h, path.encode())

    def forward(self, ids):
        """ids [B, S] -> last-position logits [B, vocab] (the graph's own output rows; TinyLLaMAModel returns all S rows)."""
        ids = np.ascontiguousarray(ids, dtype=np.float32)
        B, S = ids.shape
        out = np.empty((B * S, self.vocab), dtype=np.float32)
        so = self.L.cpuref_model_forward(self.h, _p(ids), B, S, _p(out), B * S)
        if so <= 0:
            raise RuntimeError("reference forward failed")
        return out[:B * so].reshape(B, so, self.vocab)[:, -1, :].copy()

    def generate(self, ids, max_new, end_token=-1):
        """The reference's Module::generate (greedy) on a [B, S] batch -> uint32 [B, max_new] (rows that hit end_token are shorter: n)."""
        ids = np.ascontiguousarray(ids, dtype=np.float32)
        B, S = ids.shape
        out = np.zeros((B, max_new), dtype=np.uint32)
        n = np.zeros(B, dtype=np.int32)
        self.L.cpuref_model_generate(self.h, _p(ids), B, S, max_new, end_token, _p(out), _p(n))
        return out, n

    def reset(self):
        self.L.cpuref_model_clear(self.h)

    def close(self):
        if self.h:
            self.L.cpuref_model_destroy(self.h)
            self.h = None


# ---------------------------------------------------------------------------------------------------------------- samplers
# Restatement of _LlmTextGenerateTopkSamplingMethod::generate (Generate.cpp:45-91) and _LlmTextGenerateToppSamplingMethod::generate
# (Generate.cpp:93-143) up to the random draw: the candidate ids and the probabilities handed to std::discrete_distribution, in the
# reference's float / double mix.  Pinned statistically against the reference itself (cpuref_sample_hist draws from the compiled
# classes; tests/test_sampling.py).  TEST INFRASTRUCTURE ONLY.
def _softmax_candidates(vals, temperature):
    vals = np.asarray(vals, dtype=np.float32)
    mx = np.float64(vals.max())
    e = np.exp((vals.astype(np.float64) - mx) / np.float64(np.float32(temperature))).astype(np.float32)  # softmax[i] = exp(double) -> float
    sum_exp = np.float64(0.0)
    for v in e:
        sum_exp += np.float64(v)
    p = (e.astype(np.float64) / sum_exp).astype(np.float32)                                               # value /= sum_exp (float /= double)
    s2 = np.float64(0.0)
    for v in p:
        s2 += np.float64(v)
    return (p / np.float32(s2)).astype(np.float32)                                                        # value /= _sum (float)


def sample_topk_candidates(logits, k, temperature):
    """-> (ids [k] in descending order of logit, lowest id first among equals; probabilities [k])."""
    x = np.asarray(logits, dtype=np.float32)
    order = np.lexsort((np.arange(x.size), -x.astype(np.float64)))[:k]
    return order.astype(np.int64), _softmax_candidates(x[order], temperature)


def sample_topp_candidates(probs, p, temperature):
    x = np.asarray(probs, dtype=np.float32)
    order = np.lexsort((np.arange(x.size), -x.astype(np.float64)))
    cum, n = np.float32(0.0), 0
    while cum < np.float32(p):
        cum = np.float32(cum + x[order[n]])
        n += 1
    ids = order[:n].astype(np.int64)
    if n == 1:
        return ids, np.ones(1, dtype=np.float32)
    return ids, _softmax_candidates(x[ids], temperature)


def sample_draw(probs, u):
    """std::discrete_distribution: weights / their double sum, running sums, first bucket whose running sum exceeds u."""
    w = np.asarray(probs, dtype=np.float32).astype(np.float64)
    tot = np.float64(0.0)
    for v in w:
        tot += v
    run = np.float64(0.0)
    for i, v in enumerate(w):
        run += v / tot
        if u < run:
            return i
    return len(w) - 1


def ref_sample_hist(mode, x, p_or_k, temperature, draws):
    """Counts of `draws` calls of the REFERENCE's sampler on row x (mode 0 top-k on logits, 1 top-p on probabilities)."""
    L = cpuref()
    x = np.ascontiguousarray(x, dtype=np.float32)
    counts = np.zeros(x.size, dtype=np.uint32)
    L.cpuref_sample_hist.argtypes = [C.c_int, C.c_void_p, C.c_int, C.c_float, C.c_float, C.c_int, C.c_void_p]
    rc = L.cpuref_sample_hist(mode, x.ctypes.data, x.size, C.c_float(p_or_k), C.c_float(temperature), draws, counts.ctypes.data)
    assert rc == 0
    return counts
