"""ctypes binding of the C ABI in include/mllmcu.h (libmllmcu.so).

This module is plumbing for tests and bench.py: it passes raw device pointers
(from torch tensors) and the current CUDA stream to the C entry points.  There
is no CPU fallback: a missing library raises, a non-zero return code raises.
"""
import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "lib", "libmllmcu.so")

TYPE_F32, TYPE_F16, TYPE_Q4_0, TYPE_Q8_0 = 0, 1, 2, 8
TYPE_Q4_K, TYPE_Q6_K, TYPE_Q4_0_4_4 = 12, 14, 19
Q4_K_BYTES, Q6_K_BYTES, Q8_K_BYTES = 144, 210, 292
ROPE_LLAMA, ROPE_HF = 2, 4
Q4_0_BYTES, Q8_0_BYTES = 18, 34

# every symbol include/mllmcu.h declares: (restype, argtypes)
_vp, _i, _sz, _ll, _f = C.c_void_p, C.c_int, C.c_size_t, C.c_longlong, C.c_float
SYMBOLS = {
    "mllmcu_version": (_i, []),
    "mllmcu_device_count": (_i, []),
    "mllmcu_set_device": (_i, [_i]),
    "mllmcu_device_info": (_i, [_i, C.c_char_p, C.POINTER(_i), C.POINTER(_i), C.POINTER(_i), C.POINTER(_sz)]),
    "mllmcu_malloc": (_i, [C.POINTER(_vp), _sz]),
    "mllmcu_free": (_i, [_vp]),
    "mllmcu_malloc_managed": (_i, [C.POINTER(_vp), _sz, _i]),
    "mllmcu_mem_prefetch": (_i, [_vp, _sz, _i, _vp]),
    "mllmcu_memset": (_i, [_vp, _i, _sz, _vp]),
    "mllmcu_memcpy_h2d": (_i, [_vp, _vp, _sz, _vp]),
    "mllmcu_memcpy_d2h": (_i, [_vp, _vp, _sz, _vp]),
    "mllmcu_memcpy_d2d": (_i, [_vp, _vp, _sz, _vp]),
    "mllmcu_host_alloc": (_i, [C.POINTER(_vp), _sz]),
    "mllmcu_host_free": (_i, [_vp]),
    "mllmcu_stream_create": (_i, [C.POINTER(_vp)]),
    "mllmcu_stream_destroy": (_i, [_vp]),
    "mllmcu_stream_sync": (_i, [_vp]),
    "mllmcu_event_create": (_i, [C.POINTER(_vp)]),
    "mllmcu_event_destroy": (_i, [_vp]),
    "mllmcu_event_record": (_i, [_vp, _vp]),
    "mllmcu_event_sync": (_i, [_vp]),
    "mllmcu_event_elapsed_ms": (_i, [_vp, _vp, C.POINTER(_f)]),
    "mllmcu_graph_begin_capture": (_i, [_vp]),
    "mllmcu_graph_end_capture": (_i, [_vp, C.POINTER(_vp)]),
    "mllmcu_graph_abort_capture": (_i, [_vp]),
    "mllmcu_graph_launch": (_i, [_vp, _vp]),
    "mllmcu_graph_destroy": (_i, [_vp]),
    "mllmcu_quantize_q8_0": (_i, [_vp, _vp, _i, _i, _i, _vp]),
    "mllmcu_block_isums_q4_0_q8_0": (_i, [_vp, _vp, _vp, _i, _i, _i, _vp]),
    "mllmcu_linear_workspace_bytes": (_sz, [_i, _i]),
    "mllmcu_linear": (_i, [_i, _vp, _vp, _vp, _vp, _i, _i, _i, _i, _i, _i, _vp, _sz, _vp]),
    "mllmcu_linear_ex": (_i, [_i, _vp, _vp, _vp, _vp, _i, _i, _i, _i, _i, _i, _vp, _sz, _i, _vp]),
    "mllmcu_linear_quant_kind": (_i, [_i, _i, _i, _i, _vp]),
    "mllmcu_linear_static_ws_capacity": (_sz, []),
    "mllmcu_linear_static_ws": (_i, [_i, _vp, _vp, _vp, _vp, _i, _i, _i, _i, _i, _i, _vp]),
    "mllmcu_gemv_fused": (_i, [_i, _vp, _vp, _vp, _i, _vp, _i, _i, _i, _vp, _f, _i, _vp, _i, _vp, _i, _i, _i, _vp]),
    "mllmcu_gemv_fused_ex": (_i, [_vp, _vp]),
    "mllmcu_gemv_fused_max_rows": (_i, []),
    "mllmcu_gemv_prof_read": (_i, [_vp, _i]),
    "mllmcu_decode_attn_prof_read": (_i, [_vp, _i]),
    "mllmcu_selftest_div": (_i, [C.c_uint32, C.c_uint32, _vp, _vp]),
    "mllmcu_rmsnorm": (_i, [_vp, _vp, _vp, _i, _i, _f, _i, _vp]),
    "mllmcu_layernorm": (_i, [_vp, _vp, _vp, _vp, _i, _i, _f, _vp]),
    "mllmcu_rope": (_i, [_i, _vp, _vp, _i, _i, _i, _i, _i, _i, _vp, _vp, _i, _i, _ll, _ll, _vp]),
    "mllmcu_silu": (_i, [_vp, _vp, _sz, _vp]),
    "mllmcu_silu_mul": (_i, [_vp, _vp, _vp, _sz, _vp]),
    "mllmcu_add": (_i, [_vp, _vp, _vp, _sz, _vp]),
    "mllmcu_mul": (_i, [_vp, _vp, _vp, _sz, _vp]),
    "mllmcu_embedding_q4_0": (_i, [_vp, _vp, _vp, _i, _i, _vp]),
    "mllmcu_embedding_f32": (_i, [_vp, _vp, _vp, _i, _i, _vp]),
    "mllmcu_flash_attn_workspace_bytes": (_sz, [_i, _i, _i, _i]),
    "mllmcu_flash_attn_f32kv_workspace_bytes": (_sz, [_i, _i, _i, _i, _i, _i]),
    "mllmcu_flash_attn_f16kv": (_i, [_vp, _vp, _vp, _vp, _i, _i, _i, _i, _i, _i, _i, _ll, _vp, _sz, _vp]),
    "mllmcu_quantize_q8_K": (_i, [_vp, _vp, _i, _i, _i, _vp]),
    "mllmcu_block_isums_q4_K_q8_K": (_i, [_vp, _vp, _vp, _vp, _i, _i, _i, _vp]),
    "mllmcu_prep_q8_K": (_i, [_i, _vp, _vp, C.c_float, _i, _vp, _i, _i, _i, _i, _vp]),
    "mllmcu_gemv_q4k": (_i, [_i, _vp, _vp, _vp, _vp, _i, _vp, _i, _i, _vp, _i, _i, _vp]),
    "mllmcu_q4_0_4x4_to_q4_0": (_i, [_vp, _vp, _i, _i, _vp]),
    "mllmcu_gelu_table": (_i, [_vp, _vp, _sz, _vp, _vp]),
    "mllmcu_softmax": (_i, [_vp, _vp, _i, _i, _i, _i, _i, _vp]),
    "mllmcu_div_scalar": (_i, [_vp, _f, _vp, _sz, _vp]),
    "mllmcu_rope_mrope": (_i, [_vp, _vp, _i, _i, _i, _i, _i, _vp, _vp, _i, _i, _ll, _ll, _vp]),
    "mllmcu_rope_vision": (_i, [_vp, _vp, _vp, _vp, _i, _i, _i, _i, _i, _vp]),
    "mllmcu_copy_strided": (_i, [_vp, _vp, _i, _vp, _vp, _vp, _vp]),
    "mllmcu_sgemm": (_i, [_vp, _ll, _ll, _vp, _ll, _ll, _ll, _vp, _vp, _ll, _ll, _i, _i, _i, _i, _vp]),
    "mllmcu_ipc_get_handle": (_i, [_vp, _vp]),
    "mllmcu_ipc_open": (_i, [_vp, _vp]),
    "mllmcu_ipc_close": (_i, [_vp]),
    "mllmcu_p2p_bump": (_i, [_vp, _vp]),
    "mllmcu_p2p_allreduce": (_i, [_i, _i, _vp, _sz, _sz, _vp, _vp, _vp, _i, _i, _vp]),
    "mllmcu_sample": (_i, [_i, _vp, _i, _i, _ll, _f, _f, _vp, _vp, _vp, _vp, _vp, _vp, _i, _vp]),
    "mllmcu_where_eq": (_i, [_vp, _ll, _f, _vp, _vp, _vp]),
    "mllmcu_where_unravel": (_i, [_vp, _i, _vp, _i, _vp, _vp]),
    "mllmcu_index_put_rows": (_i, [_vp, _ll, _vp, _ll, _vp, _i, _i, _vp]),
    "mllmcu_index_splice": (_i, [_vp, _vp, _vp, _vp, _i, _i, _i, _vp]),
    "mllmcu_range": (_i, [_vp, _i, _i, _vp]),
    "mllmcu_conv2d_valid": (_i, [_vp, _vp, _vp, _vp, _i, _i, _i, _i, _i, _i, _i, _i, _i, _i, _vp]),
    "mllmcu_conv3d_valid": (_i, [_vp, _vp, _vp, _vp, _i, _i, _i, _i, _i, _i, _i, _i, _i, _i, _i, _i, _i, _vp]),
    "mllmcu_flash_attn_f32kv": (_i, [_vp, _vp, _vp, _vp, _i, _i, _i, _i, _i, _i, _i, _ll, _vp, _sz, _vp]),
    "mllmcu_f32_to_f16_rows": (_i, [_vp, _vp, _i, _i, _ll, _ll, _vp]),
    "mllmcu_f16_to_f32_rows": (_i, [_vp, _vp, _i, _i, _ll, _ll, _vp]),
    "mllmcu_argmax": (_i, [_vp, _i, _i, _vp, _vp]),
    "mllmcu_argmax_ids": (_i, [_vp, _i, _i, _vp, _vp]),
    "mllmcu_decode_embed": (_i, [_vp, _vp, _vp, _i, _i, _vp, _vp]),
    "mllmcu_decode_rope_kv": (_i, [_i, _vp, _i, _i, _i, _i, _i, _i, _vp, _vp, _i, _vp, _vp, _vp, C.c_longlong, _i, _i, _vp]),
    "mllmcu_decode_attn_nsplit": (_i, [_i, _i]),
    "mllmcu_decode_attn_workspace_bytes": (_sz, [_i, _i, _i]),
    "mllmcu_decode_attn": (_i, [_vp, _i, _vp, _vp, _vp, _i, _vp, _i, _i, _i, _i, C.c_longlong, _vp, _sz, _i, _vp]),
    "mllmcu_decode_attn_rope": (_i, [_i, _vp, _i, _vp, _vp, _vp, _i, _vp, _i, _i, _i, _i, _vp, _vp, _i, C.c_longlong, _i, _vp, _sz, _i, _vp]),
    "mllmcu_decode_argmax": (_i, [_vp, _i, _i, _vp, _vp, _vp, _i, _vp]),
    "mllmcu_decode_step_create": (_i, [_vp, _vp, _i, C.POINTER(_vp)]),
    "mllmcu_decode_step_launch": (_i, [_vp, _vp]),
    "mllmcu_decode_step_destroy": (_i, [_vp]),
}

_lib = None


class MllmcuError(RuntimeError):
    pass


def lib():
    """Load libmllmcu.so (raises if it has not been built -- there is no fallback)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise MllmcuError(f"{LIB_PATH} missing: run `python -m mllm_b200.build` (no CPU fallback exists)")
        L = C.CDLL(LIB_PATH)
        for name, (res, args) in SYMBOLS.items():
            fn = getattr(L, name)  # AttributeError if the library does not export a declared symbol
            fn.restype = res
            fn.argtypes = args
        _lib = L
    return _lib


def check(rc, what=""):
    if rc != 0:
        raise MllmcuError(f"{what} failed with code {rc}")


def _stream():
    import torch
    return torch.cuda.current_stream().cuda_stream


def _ptr(t):
    return None if t is None else t.data_ptr()


# ---------------------------------------------------------------- torch-facing helpers
def quantize_q8_0(x):
    """x: cuda fp32 [rows,k] contiguous -> cuda uint8 [rows, k/32*34] native block_q8_0."""
    import torch
    rows, k = x.shape
    y = torch.empty((rows, k // 32 * Q8_0_BYTES), dtype=torch.uint8, device=x.device)
    check(lib().mllmcu_quantize_q8_0(_ptr(x), _ptr(y), rows, k, x.stride(0), _stream()), "quantize_q8_0")
    return y


def block_isums_q4_0_q8_0(W, xq, K):
    import torch
    N, M = W.shape[0], xq.shape[0]
    out = torch.empty((N, M, K // 32), dtype=torch.int32, device=W.device)
    check(lib().mllmcu_block_isums_q4_0_q8_0(_ptr(W), _ptr(xq), _ptr(out), M, N, K, _stream()), "block_isums")
    return out


_ws_cache = {}


def _workspace(nbytes, device):
    import torch
    key = (device.index, "ws")
    t = _ws_cache.get(key)
    if t is None or t.numel() < nbytes:
        t = torch.empty(max(nbytes, 1 << 20), dtype=torch.uint8, device=device)
        _ws_cache[key] = t
    return t


def linear(W, x, bias=None, wtype=TYPE_Q4_0, out_f16=False, out=None, ldy=None, reuse_quant=False):
    """x cuda fp32 [M,K]; W cuda uint8 [N, K/32*blk]; returns y [M,N] fp32 (or fp16).
    reuse_quant: the workspace still holds the quantised image of x from the previous call (mllmcu_linear_ex flag 1)."""
    import torch
    M, K = x.shape
    N = W.shape[0]
    if out is None:
        out = torch.empty((M, N), dtype=torch.float16 if out_f16 else torch.float32, device=x.device)
        ldy = N
    need = lib().mllmcu_linear_workspace_bytes(M, K)
    ws = _workspace(need, x.device)
    if reuse_quant and lib().mllmcu_linear_quant_kind(wtype, M, N, K, _ptr(W)) == 0:
        raise MllmcuError("linear: this shape keeps no reusable quantised image")
    check(lib().mllmcu_linear_ex(wtype, _ptr(W), _ptr(x), _ptr(bias), _ptr(out), M, N, K, x.stride(0), ldy, int(out_f16),
                                 _ptr(ws), ws.numel(), int(reuse_quant), _stream()), "linear")
    return out


def linear_static_ws(W, x, bias=None, wtype=TYPE_Q4_0, out_f16=False):
    """mllmcu_linear_static_ws: the Linear with the library's own static device scratch (the call allocates nothing;
    the output tensor is made by the caller -- here -- beforehand).  Returns (rc, y): rc is NOT raised so tests can see -3."""
    import torch
    M, K = x.shape
    N = W.shape[0]
    out = torch.empty((M, N), dtype=torch.float16 if out_f16 else torch.float32, device=x.device)
    rc = lib().mllmcu_linear_static_ws(wtype, _ptr(W), _ptr(x), _ptr(bias), _ptr(out), M, N, K, x.stride(0), N, int(out_f16), _stream())
    return rc, out


def rmsnorm(x, w, eps, add_unit_offset=False):
    import torch
    y = torch.empty_like(x)
    rows = x.numel() // x.shape[-1]
    check(lib().mllmcu_rmsnorm(_ptr(x), _ptr(w), _ptr(y), rows, x.shape[-1], eps, int(add_unit_offset), _stream()), "rmsnorm")
    return y


def layernorm(x, w, b, eps):
    import torch
    y = torch.empty_like(x)
    rows = x.numel() // x.shape[-1]
    check(lib().mllmcu_layernorm(_ptr(x), _ptr(w), _ptr(b), _ptr(y), rows, x.shape[-1], eps, _stream()), "layernorm")
    return y


def rope(rope_type, x, pos0, sin_t, cos_t, rot=None, out_f16=False, out=None, out_bstride=None, out_sstride=None):
    """x cuda fp32 [B,S,H,D]."""
    import torch
    B, S, H, D = x.shape
    rot = D if rot is None else rot
    if out is None:
        out = torch.empty((B, S, H, D), dtype=torch.float16 if out_f16 else torch.float32, device=x.device)
        out_bstride, out_sstride = S * H * D, H * D
    check(lib().mllmcu_rope(rope_type, _ptr(x), _ptr(out), B, S, H, D, rot, pos0, _ptr(sin_t), _ptr(cos_t), sin_t.shape[1],
                            int(out_f16), out_bstride, out_sstride, _stream()), "rope")
    return out


def _ew(fn, a, b=None):
    import torch
    y = torch.empty_like(a)
    if b is None:
        check(fn(_ptr(a), _ptr(y), a.numel(), _stream()))
    else:
        check(fn(_ptr(a), _ptr(b), _ptr(y), a.numel(), _stream()))
    return y


def silu(x):
    return _ew(lib().mllmcu_silu, x)


def silu_mul(g, u):
    return _ew(lib().mllmcu_silu_mul, g, u)


def add(a, b):
    return _ew(lib().mllmcu_add, a, b)


def mul(a, b):
    return _ew(lib().mllmcu_mul, a, b)


def embedding_q4_0(table, ids, hidden):
    import torch
    n = ids.numel()
    out = torch.zeros((n, hidden), dtype=torch.float32, device=table.device)
    check(lib().mllmcu_embedding_q4_0(_ptr(table), _ptr(ids), _ptr(out), n, hidden, _stream()), "embedding")
    return out


def flash_attn_f16kv(q, k, v, causal=True, kv_bstride=None):
    """q cuda fp32 [B,Sq,Hq,D]; k,v cuda fp16 [B,Sk,Hkv,D] (or a cache with batch stride kv_bstride)."""
    import torch
    B, Sq, Hq, D = q.shape
    Sk, Hkv = k.shape[1], k.shape[2]
    kv_bstride = k.stride(0) if kv_bstride is None else kv_bstride
    o = torch.empty_like(q)
    need = lib().mllmcu_flash_attn_workspace_bytes(B, Sq, Hq, D)
    ws = _workspace(need, q.device)
    check(lib().mllmcu_flash_attn_f16kv(_ptr(q), _ptr(k), _ptr(v), _ptr(o), B, Sq, Sk, Hq, Hkv, D, int(causal), kv_bstride,
                                        _ptr(ws), ws.numel(), _stream()), "flash_attn")
    return o


# ---------------------------------------------------------------- rows a4/a5/a7/a10/a11/a15/a17/a18
def quantize_q8_K(x):
    """x: cuda fp32 [rows,k] -> cuda uint8 [rows, k/256*292] native block_q8_K."""
    import torch
    rows, k = x.shape
    y = torch.zeros((rows, k // 256 * Q8_K_BYTES), dtype=torch.uint8, device=x.device)
    check(lib().mllmcu_quantize_q8_K(_ptr(x), _ptr(y), rows, k, x.stride(0), _stream()), "quantize_q8_K")
    return y


Q8_K_PITCH = 304  # the decode chain's re-pitched block_q8_K: d at 0, qs at 16, the 16 group sums at 272 (bytes 4..15 unused)


def prep_q8_K(x, mode=0, aux=None, eps=0.0, add_unit_offset=False, pdl=False, aligned=False):
    """block_q8_K rows of x (mode 0), RMSNorm(x; aux) (mode 1) or SiLU(x) * aux (mode 2): x cuda fp32 [rows,k].
    aligned: 304-byte blocks (see Q8_K_PITCH) instead of the native 292-byte ones."""
    import torch
    rows, k = x.shape
    y = torch.zeros((rows, k // 256 * (Q8_K_PITCH if aligned else Q8_K_BYTES)), dtype=torch.uint8, device=x.device)
    check(lib().mllmcu_prep_q8_K(mode, _ptr(x), _ptr(aux), float(eps), int(add_unit_offset), _ptr(y), rows, k,
                                 x.stride(0), int(pdl) | (2 if aligned else 0), _stream()), "prep_q8_K")
    return y


def q8_K_unpitch(y):
    """304-byte blocks -> native 292-byte block_q8_K bytes."""
    import torch
    b = y.reshape(y.shape[0], -1, Q8_K_PITCH)
    return torch.cat([b[:, :, 0:4], b[:, :, 16:304]], dim=2).reshape(y.shape[0], -1).contiguous()


def gemv_q4k(Ws, xq, K, biases=None, residual=None, pdl=False, aligned=False, no_areg=False):
    """Ws: 1..3 cuda uint8 [N_s, K/256*144] Q4_K matrices; xq: block_q8_K rows [M, ...] (native, or 304-byte blocks with aligned=True)
    -> list of fp32 [M, N_s] (one launch)."""
    import torch
    M, n = xq.shape[0], len(Ws)
    ys = [torch.empty((M, W.shape[0]), dtype=torch.float32, device=xq.device) for W in Ws]
    ld = ys[0].stride(0)
    if n > 1:  # one leading dimension for every segment: views of one [M, sum N] buffer
        big = torch.empty((M, sum(W.shape[0] for W in Ws)), dtype=torch.float32, device=xq.device)
        off, ys = 0, []
        for W in Ws:
            ys.append(big[:, off:off + W.shape[0]])
            off += W.shape[0]
        ld = big.stride(0)
    Wp = (C.c_void_p * n)(*[_ptr(W) for W in Ws])
    Np = (C.c_int * n)(*[int(W.shape[0]) for W in Ws])
    bp = (C.c_void_p * n)(*[(_ptr(b) if b is not None else None) for b in (biases or [None] * n)])
    yp = (C.c_void_p * n)(*[y.data_ptr() for y in ys])
    check(lib().mllmcu_gemv_q4k(n, Wp, Np, bp, yp, ld, _ptr(xq), M, K, _ptr(residual) if residual is not None else None,
                                residual.stride(0) if residual is not None else 0, int(pdl) | (2 if aligned else 0) | (4 if no_areg else 0), _stream()), "gemv_q4k")
    return ys


def block_isums_q4_K_q8_K(W, xq, K):
    import torch
    N, M = W.shape[0], xq.shape[0]
    a = torch.empty((N, M, K // 256), dtype=torch.int32, device=W.device)
    b = torch.empty_like(a)
    check(lib().mllmcu_block_isums_q4_K_q8_K(_ptr(W), _ptr(xq), _ptr(a), _ptr(b), M, N, K, _stream()), "block_isums_q4_K")
    return a, b


def q4_0_4x4_to_q4_0(packed, nrow, k):
    import torch
    out = torch.empty((nrow, k // 32 * Q4_0_BYTES), dtype=torch.uint8, device=packed.device)
    check(lib().mllmcu_q4_0_4x4_to_q4_0(_ptr(packed), _ptr(out), nrow, k, _stream()), "q4_0_4x4_to_q4_0")
    return out


def gelu_table(x, table):
    import torch
    y = torch.empty_like(x)
    check(lib().mllmcu_gelu_table(_ptr(x), _ptr(y), x.numel(), _ptr(table), _stream()), "gelu_table")
    return y


def softmax(x, causal=False, old_dim=0):
    """x cuda fp32 [outer, S, n]"""
    import torch
    outer, S, n = x.shape
    y = torch.empty_like(x)
    check(lib().mllmcu_softmax(_ptr(x), _ptr(y), outer, S, n, int(causal), old_dim, _stream()), "softmax")
    return y


def div_scalar(x, s):
    import torch
    y = torch.empty_like(x)
    check(lib().mllmcu_div_scalar(_ptr(x), s, _ptr(y), x.numel(), _stream()), "div_scalar")
    return y


def rope_mrope(x, sin_t, cos_t, out_f16=False):
    import torch
    B, S, H, D = x.shape
    out = torch.empty((B, S, H, D), dtype=torch.float16 if out_f16 else torch.float32, device=x.device)
    check(lib().mllmcu_rope_mrope(_ptr(x), _ptr(out), B, S, H, D, D, _ptr(sin_t), _ptr(cos_t), sin_t.shape[1], int(out_f16),
                                  S * H * D, H * D, _stream()), "rope_mrope")
    return out


def rope_vision(x, angles=None, sin_t=None, cos_t=None):
    import torch
    B, S, H, D = x.shape
    out = torch.empty_like(x)
    a0 = angles if angles is not None else sin_t
    check(lib().mllmcu_rope_vision(_ptr(x), _ptr(a0), _ptr(cos_t), _ptr(out), B, S, H, D, 0, _stream()), "rope_vision")
    return out


def copy_strided(src, dst, n, ss, ds):
    n_ = (C.c_int * 4)(*n)
    ss_ = (C.c_longlong * 4)(*ss)
    ds_ = (C.c_longlong * 4)(*ds)
    check(lib().mllmcu_copy_strided(_ptr(src), _ptr(dst), src.element_size(), n_, ss_, ds_, _stream()), "copy_strided")
    return dst


def sgemm(A, B, bias=None, b_is_kn=False):
    """A cuda fp32 [batch, M, K]; B [batch, N, K] (or [batch, K, N] with b_is_kn) -> C [batch, M, N]"""
    import torch
    batch, M, K = A.shape
    N = B.shape[2] if b_is_kn else B.shape[1]
    Cm = torch.empty((batch, M, N), dtype=torch.float32, device=A.device)
    ldbn, ldbk = (1, N) if b_is_kn else (K, 1)
    check(lib().mllmcu_sgemm(_ptr(A), K, M * K, _ptr(B), ldbn, ldbk, B.stride(0), _ptr(bias), _ptr(Cm), N, M * N, M, N, K, batch, _stream()), "sgemm")
    return Cm


def conv2d_valid(x, w, bias, stride, channels_last=False):
    import torch
    B, Cc, H, W = x.shape
    O, _, kh, kw = w.shape
    oh, ow = (H - kh) // stride[0] + 1, (W - kw) // stride[1] + 1
    y = torch.empty((B, oh * ow, O) if channels_last else (B, O, oh, ow), dtype=torch.float32, device=x.device)
    check(lib().mllmcu_conv2d_valid(_ptr(x), _ptr(w), _ptr(bias), _ptr(y), B, Cc, H, W, O, kh, kw, stride[0], stride[1], int(channels_last), _stream()), "conv2d")
    return y


def conv3d_valid(x, w, bias, stride, channels_last=False):
    import torch
    B, Cc, T, H, W = x.shape
    O, _, kt, kh, kw = w.shape
    ot, oh, ow = (T - kt) // stride[0] + 1, (H - kh) // stride[1] + 1, (W - kw) // stride[2] + 1
    y = torch.empty((B, ot, oh, ow, O) if channels_last else (B, O, ot, oh, ow), dtype=torch.float32, device=x.device)
    check(lib().mllmcu_conv3d_valid(_ptr(x), _ptr(w), _ptr(bias), _ptr(y), B, Cc, T, H, W, O, kt, kh, kw, stride[0], stride[1], stride[2],
                                    int(channels_last), _stream()), "conv3d")
    return y


def flash_attn_f32kv(q, k, v, causal=False, small_ws=False):
    """q cuda fp32 [B,Sq,Hq,D]; k,v cuda fp32 [B,Sk,Hkv,D].  small_ws: only the split-K scratch (K/V are then split tile by tile in the kernel)."""
    import torch
    B, Sq, Hq, D = q.shape
    Sk, Hkv = k.shape[1], k.shape[2]
    o = torch.empty_like(q)
    need = lib().mllmcu_flash_attn_workspace_bytes(B, Sq, Hq, D) if small_ws else lib().mllmcu_flash_attn_f32kv_workspace_bytes(B, Sq, Sk, Hq, Hkv, D)
    ws = _workspace(need, q.device)
    check(lib().mllmcu_flash_attn_f32kv(_ptr(q), _ptr(k), _ptr(v), _ptr(o), B, Sq, Sk, Hq, Hkv, D, int(causal), k.stride(0),
                                        _ptr(ws), need, _stream()), "flash_attn_f32kv")
    return o


def argmax(logits):
    import torch
    rows, n = logits.shape
    out = torch.empty((rows,), dtype=torch.int32, device=logits.device)
    check(lib().mllmcu_argmax(_ptr(logits), rows, n, _ptr(out), _stream()), "argmax")
    return out


class GemvDesc(C.Structure):
    """mllmcu_gemv_desc (include/mllmcu.h)"""
    _fields_ = [("wtype", C.c_int), ("nW", C.c_int), ("W", C.c_void_p * 3), ("N", C.c_int * 3), ("bias", C.c_void_p * 3), ("x", C.c_void_p),
                ("ldx", C.c_int), ("M", C.c_int), ("K", C.c_int), ("norm_w", C.c_void_p), ("eps", C.c_float), ("epilogue", C.c_int),
                ("residual", C.c_void_p), ("ldr", C.c_int), ("y", C.c_void_p), ("ldy", C.c_int), ("out_f16", C.c_int), ("pdl", C.c_int),
                ("x_reps", C.c_int), ("norm_reps", C.c_int), ("y_reps", C.c_int), ("x_rep_stride", C.c_longlong),
                ("norm_rep_stride", C.c_longlong), ("y_rep_stride", C.c_longlong), ("n_pf", C.c_int), ("pf_ptr", C.c_void_p * 8),
                ("pf_bytes", C.c_size_t * 8), ("xchg_world", C.c_int), ("xchg_rank", C.c_int), ("xchg_peer_bases", C.c_void_p * 8),
                ("xchg_buf_offset", C.c_size_t), ("xchg_flag_offset", C.c_size_t), ("xchg_epoch", C.c_void_p), ("xchg_ticket", C.c_void_p),
                ("xchg_residual", C.c_void_p), ("xchg_out", C.c_void_p)]


def gemv_fused_ex(Ws, x, y, norm_w=None, eps=1e-6, epilogue=0, residual=None, pdl=False, x_reps=1, norm_reps=1, y_reps=1, lookahead=(), xchg=None):
    """Descriptor form of gemv_fused.  x: cuda fp32 [x_reps, M, K]; norm_w: [norm_reps, K]; y: preallocated [y_reps, M, N];
    residual: [M, N] (copy 0); lookahead: (device pointer, bytes) pairs for the L2 look-ahead."""
    d = GemvDesc()
    d.wtype, d.nW = TYPE_Q4_0, len(Ws)
    for i, w in enumerate(Ws):
        d.W[i], d.N[i] = w.data_ptr(), w.shape[0]
    M, K = x.shape[-2], x.shape[-1]
    d.x, d.ldx, d.M, d.K = x.data_ptr(), K, M, K
    if norm_w is not None:
        d.norm_w, d.eps, d.norm_reps, d.norm_rep_stride = norm_w.data_ptr(), eps, norm_reps, K
    d.epilogue = epilogue
    if residual is not None:
        d.residual, d.ldr = residual.data_ptr(), residual.stride(-2)
    NT = y.shape[-1]
    d.y, d.ldy, d.out_f16, d.pdl = y.data_ptr(), NT, int(y.element_size() == 2), int(pdl)
    d.x_reps, d.x_rep_stride, d.y_reps, d.y_rep_stride = x_reps, M * K, y_reps, M * NT
    d.n_pf = len(lookahead)
    for i, (ptr, nbytes) in enumerate(lookahead):
        d.pf_ptr[i], d.pf_bytes[i] = ptr, nbytes
    if xchg is not None:  # dict(world, rank, bases (c_void_p * world), buf_offset, flag_offset, epoch, ticket, residual, out): see mllmcu_gemv_desc.xchg_*
        d.xchg_world, d.xchg_rank = xchg["world"], xchg["rank"]
        for r in range(xchg["world"]):
            d.xchg_peer_bases[r] = xchg["bases"][r]
        d.xchg_buf_offset, d.xchg_flag_offset = xchg["buf_offset"], xchg["flag_offset"]
        d.xchg_epoch, d.xchg_ticket = xchg["epoch"].data_ptr(), xchg["ticket"].data_ptr()
        d.xchg_residual = xchg["residual"].data_ptr() if xchg.get("residual") is not None else None
        d.xchg_out = xchg["out"].data_ptr()
    check(lib().mllmcu_gemv_fused_ex(C.byref(d), _stream()), "gemv_fused_ex")
    return y


def gemv_fused(Ws, x, norm_w=None, eps=1e-6, biases=None, epilogue=0, residual=None, out_f16=False, pdl=False):
    """Fused decode Linear; Ws: list of cuda uint8 [N_i, K/32*18]; x cuda fp32 [M,K]. Returns y [M, sum N] (or [M, N0] for epilogue 2)."""
    import torch
    M, K = x.shape
    nW = len(Ws)
    Wp = (C.c_void_p * nW)(*[w.data_ptr() for w in Ws])
    Ns = (C.c_int * nW)(*[w.shape[0] for w in Ws])
    bp = None
    if biases is not None:
        bp = (C.c_void_p * nW)(*[(b.data_ptr() if b is not None else None) for b in biases])
    NT = Ws[0].shape[0] if epilogue == 2 else sum(w.shape[0] for w in Ws)
    y = torch.empty((M, NT), dtype=torch.float16 if out_f16 else torch.float32, device=x.device)
    rc = lib().mllmcu_gemv_fused(TYPE_Q4_0, Wp, Ns, bp, nW, _ptr(x), x.stride(0), M, K, _ptr(norm_w), eps, epilogue,
                                 _ptr(residual), residual.stride(0) if residual is not None else 0, _ptr(y), NT, int(out_f16), int(pdl), _stream())
    check(rc, "gemv_fused")
    return y


# ---------------------------------------------------------------- decode-step kernels (graph-replayable)
def decode_embed(table, ids, hidden, ctl=None):
    import torch
    n = ids.numel()
    out = torch.empty((n, hidden), dtype=torch.float32, device=table.device)
    check(lib().mllmcu_decode_embed(_ptr(table), _ptr(ids), _ptr(out), n, hidden, _ptr(ctl), _stream()), "decode_embed")
    return out


def decode_rope_kv(rope_type, qkv, Hq, Hkv, D, sin_t, cos_t, ctl, kcache, vcache, rot=None, pdl=False):
    """qkv cuda fp32 [B, (Hq+2Hkv)*D] (q rotated in place); caches cuda fp16 [B, limit, Hkv, D]; ctl cuda int32 [>=1]."""
    B = qkv.shape[0]
    rot = D if rot is None else rot
    check(lib().mllmcu_decode_rope_kv(rope_type, _ptr(qkv), qkv.stride(0), B, Hq, Hkv, D, rot, _ptr(sin_t), _ptr(cos_t), sin_t.shape[1],
                                      _ptr(ctl), _ptr(kcache), _ptr(vcache), kcache.stride(0), kcache.shape[1], int(pdl), _stream()), "decode_rope_kv")


_attn_ws = {}


def decode_attn(q, ldq, kcache, vcache, ctl, Hq, D, pdl=False):
    """q cuda fp32 rows (head h at column h*D, row stride ldq); caches fp16 [B, limit, Hkv, D]; KV length = ctl[0]+1. Returns [B, Hq*D]."""
    import torch
    B, Hkv = kcache.shape[0], kcache.shape[2]
    need = lib().mllmcu_decode_attn_workspace_bytes(B, Hq, D)
    key = (q.device.index, B, Hq, D)
    ws = _attn_ws.get(key)
    if ws is None:
        ws = torch.zeros(need, dtype=torch.uint8, device=q.device)
        _attn_ws[key] = ws
    o = torch.empty((B, Hq * D), dtype=torch.float32, device=q.device)
    check(lib().mllmcu_decode_attn(_ptr(q), ldq, _ptr(kcache), _ptr(vcache), _ptr(o), Hq * D, _ptr(ctl), B, Hq, Hkv, D, kcache.stride(0),
                                   _ptr(ws), ws.numel(), int(pdl), _stream()), "decode_attn")
    return o


def decode_attn_rope(rope_type, qkv, kcache, vcache, ctl, Hq, D, sin_t, cos_t, pdl=False):
    """Fused q/k RoPE + KV append + decode attention on raw q|k|v rows [B, (Hq+2Hkv)*D]. Returns [B, Hq*D]."""
    import torch
    B, Hkv = kcache.shape[0], kcache.shape[2]
    need = lib().mllmcu_decode_attn_workspace_bytes(B, Hq, D)
    key = (qkv.device.index, B, Hq, D)
    ws = _attn_ws.get(key)
    if ws is None:
        ws = torch.zeros(need, dtype=torch.uint8, device=qkv.device)
        _attn_ws[key] = ws
    o = torch.empty((B, Hq * D), dtype=torch.float32, device=qkv.device)
    check(lib().mllmcu_decode_attn_rope(rope_type, _ptr(qkv), qkv.stride(0), _ptr(kcache), _ptr(vcache), _ptr(o), Hq * D, _ptr(ctl), B, Hq, Hkv, D,
                                        _ptr(sin_t), _ptr(cos_t), sin_t.shape[1], kcache.stride(0), kcache.shape[1], _ptr(ws), ws.numel(), int(pdl),
                                        _stream()), "decode_attn_rope")
    return o


def sample(x, u, k=None, p=None, temperature=1.0, cand_cap=64):
    """Top-k (k given; x = logits) or top-p (p given; x = probabilities) sampling of each row with the uniform numbers u [rows].
    Returns (ids fp32 [rows], candidate probabilities [rows, cand_cap], candidate ids int32 [rows, cand_cap], candidate counts int32 [rows])."""
    import torch
    rows, n = x.shape
    ws = torch.empty((rows, n), dtype=torch.float32, device=x.device)
    ids = torch.empty((rows,), dtype=torch.float32, device=x.device)
    cp = torch.zeros((rows, cand_cap), dtype=torch.float32, device=x.device)
    ci = torch.full((rows, cand_cap), -1, dtype=torch.int32, device=x.device)
    nc = torch.zeros((rows,), dtype=torch.int32, device=x.device)
    check(lib().mllmcu_sample(0 if k is not None else 1, _ptr(x), rows, n, x.stride(0), float(k if k is not None else p), float(temperature), _ptr(u), _ptr(ws), _ptr(ids),
                              _ptr(cp), _ptr(ci), _ptr(nc), cand_cap, _stream()), "sample")
    return ids, cp, ci, nc


def decode_argmax(logits, ctl=None, tok_log=None):
    import torch
    rows, n = logits.shape
    ids = torch.empty((rows,), dtype=torch.float32, device=logits.device)
    check(lib().mllmcu_decode_argmax(_ptr(logits), rows, n, _ptr(ids), _ptr(tok_log), _ptr(ctl), tok_log.shape[0] if tok_log is not None else 0,
                                     _stream()), "decode_argmax")
    return ids
