"""ctypes front-end of libmllm_cuda_backend.so: the host-side CUDA backend (CUDABackend, ops,
ParamLoader, LLaMA graph) driven with HOST numpy buffers, the way an mllm demo drives a Module.
No CPU fallback: a missing library or device raises."""
import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "lib", "libmllm_cuda_backend.so")
_lib = None


class BackendError(RuntimeError):
    pass


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise BackendError(f"{LIB_PATH} missing: run `python -m mllm_b200.build`")
        C.CDLL(os.path.join(HERE, "lib", "libmllmcu.so"), mode=C.RTLD_GLOBAL)
        L = C.CDLL(LIB_PATH)
        L.mllm_cuda_last_error.restype = C.c_char_p
        L.mllm_cuda_llama_last_forward_ms.restype = C.c_double
        L.mllm_cuda_llama_last_forward_ms.argtypes = [C.c_void_p]
        L.mllm_cuda_kernel_launches.restype = C.c_longlong
        L.mllm_cuda_llama_create.argtypes = [C.c_void_p, C.c_void_p, C.POINTER(C.c_void_p)]
        L.mllm_cuda_llama_destroy.argtypes = [C.c_void_p]
        L.mllm_cuda_llama_load.argtypes = [C.c_void_p, C.c_char_p]
        L.mllm_cuda_llama_clear_kvcache.argtypes = [C.c_void_p]
        L.mllm_cuda_llama_forward.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p]
        L.mllm_cuda_llama_block_forward.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_void_p]
        L.mllm_cuda_llama_generate.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p]
        L.mllm_cuda_stream.restype = C.c_void_p
        L.mllm_cuda_llama_last_step_fused.argtypes = [C.c_void_p]
        L.mllm_cuda_llama_set_eager_decode.argtypes = [C.c_void_p, C.c_int]
        L.mllm_cuda_llama_decode_mode.argtypes = [C.c_void_p]
        L.mllm_cuda_llama_set_step_kernel.argtypes = [C.c_void_p, C.c_int]
        L.mllm_cuda_llama_request_device.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.POINTER(C.c_float),
                                                     C.POINTER(C.c_float), C.c_void_p]
        L.mllm_cuda_convert_fp_roundtrip.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p]
        L.mllm_cuda_run_op.argtypes = [C.POINTER(C.c_char_p), C.c_void_p, C.c_int, C.c_char_p, C.c_char_p, C.POINTER(C.c_void_p),
                                       C.c_void_p, C.c_int, C.c_void_p, C.c_void_p]
        _lib = L
    return _lib


def _check(rc):
    if rc != 0:
        raise BackendError(lib().mllm_cuda_last_error().decode())


def init_backend(device=0):
    """Module::initBackend `case MLLM_CUDA` -- Backend::global_backends[MLLM_CUDA]."""
    _check(lib().mllm_cuda_init_backend(device))


def convert_fp_roundtrip(x):
    """fp32 [rows, cols] -> (fp16 bits, fp32 again) through Backend::convert_fp_data (Tensor::half() / Tensor::fp32())."""
    init_backend()
    x = np.ascontiguousarray(x, dtype=np.float32)
    h = np.empty(x.shape, dtype=np.uint16)
    b = np.empty(x.shape, dtype=np.float32)
    _check(lib().mllm_cuda_convert_fp_roundtrip(_p(x), x.shape[0], x.shape[1], _p(h), _p(b)))
    return h, b


def kernel_launches():
    return int(lib().mllm_cuda_kernel_launches())


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


class LlamaModel:
    """model = LlamaModel(cfg); model.load(path); logits = model(ids)  (examples/demo_llama.cpp shape)."""

    def __init__(self, cfg, device=0):
        init_backend(device)
        self.cfg = cfg
        ci = np.array([cfg["vocab"], cfg["hidden"], cfg["layers"], cfg["heads"], cfg["kv_heads"], cfg["ffn"], cfg["rope_type"],
                       cfg["cache_limit"], cfg["pos_max"], cfg["qkv_bias"]], dtype=np.int32)
        cf = np.array([cfg["rope_theta"], cfg["rms_eps"]], dtype=np.float32)
        self.h = C.c_void_p()
        _check(lib().mllm_cuda_llama_create(_p(ci), _p(cf), C.byref(self.h)))

    def load(self, path):
        _check(lib().mllm_cuda_llama_load(self.h, path.encode()))
        return self

    def clear_kvcache(self):
        _check(lib().mllm_cuda_llama_clear_kvcache(self.h))

    def forward(self, ids, dump=False):
        ids = np.ascontiguousarray(ids, dtype=np.float32)
        B, S = ids.shape
        logits = np.empty((B, self.cfg["vocab"]), dtype=np.float32)
        d = np.empty((self.cfg["layers"] + 1, B * S, self.cfg["hidden"]), dtype=np.float32) if dump else None
        _check(lib().mllm_cuda_llama_forward(self.h, _p(ids), B, S, _p(logits), _p(d) if dump else None))
        return (logits, d) if dump else logits

    __call__ = forward

    def block_forward(self, layer, x):
        """Block `layer` alone on the hidden state x [B, S, hidden] with an empty KV cache (teacher-forced parity tests)."""
        x = np.ascontiguousarray(x, dtype=np.float32)
        B, S, _ = x.shape
        y = np.empty_like(x)
        _check(lib().mllm_cuda_llama_block_forward(self.h, layer, _p(x), B, S, _p(y)))
        return y

    def generate(self, ids, max_new):
        ids = np.ascontiguousarray(ids, dtype=np.float32)
        B, S = ids.shape
        out = np.empty((B, max_new), dtype=np.int32)
        _check(lib().mllm_cuda_llama_generate(self.h, _p(ids), B, S, max_new, _p(out)))
        return out

    def request_device(self, ids, n_decode):
        """Prefill + n_decode greedy steps with inputs resident in HBM; returns (ms_prefill, ms_decode, tokens[B,n_decode])
        timed with CUDA events on the backend stream."""
        ids = np.ascontiguousarray(ids, dtype=np.float32)
        B, S = ids.shape
        toks = np.zeros((B, max(n_decode, 1)), dtype=np.int32)
        a, b = C.c_float(0), C.c_float(0)
        _check(lib().mllm_cuda_llama_request_device(self.h, _p(ids), B, S, n_decode, C.byref(a), C.byref(b), _p(toks)))
        return a.value, b.value, toks[:, :n_decode]

    def last_step_fused(self):
        """True when the last forward ran as the fused, graph-replayed decode step."""
        return bool(lib().mllm_cuda_llama_last_step_fused(self.h))

    def decode_mode(self):
        """0 no plan, 1 chained fused kernels, 2 persistent step kernel; +4 replayed from a CUDA graph, +8 PDL edges."""
        return int(lib().mllm_cuda_llama_decode_mode(self.h))

    def set_step_kernel(self, on):
        """Opt in to the persistent whole-step decode kernel (decode_step.cu) instead of the chained fused kernels."""
        _check(lib().mllm_cuda_llama_set_step_kernel(self.h, int(bool(on))))

    def set_eager_decode(self, on):
        """Force the op-by-op path for S == 1 (parity tests compare it with the fused step)."""
        _check(lib().mllm_cuda_llama_set_eager_decode(self.h, int(bool(on))))

    def last_forward_ms(self):
        return float(lib().mllm_cuda_llama_last_forward_ms(self.h))

    def close(self):
        if self.h:
            lib().mllm_cuda_llama_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def run_op(params, name, inputs, path="", shapes=None):
    """One op through Backend::opCreate + runOp (the test/cpu/CPUTest.hpp per-op pattern).
    params: dict OpParam (must hold 'type'); inputs: list of fp32 arrays shaped [b,h,s,d] -- or, with `shapes`
    (a list of logical (b,h,s,d) per input), arrays already in BSHD memory order."""
    init_backend()
    keys = (C.c_char_p * len(params))(*[k.encode() for k in params])
    vals = np.array(list(params.values()), dtype=np.float32)
    ins = [np.ascontiguousarray(a, dtype=np.float32) for a in inputs]
    ptrs = (C.c_void_p * len(ins))(*[a.ctypes.data for a in ins])
    shapes = np.array([a.shape for a in ins] if shapes is None else shapes, dtype=np.int32)
    out = np.empty(1 << 24, dtype=np.float32)
    oshape = np.zeros(4, dtype=np.int32)
    _check(lib().mllm_cuda_run_op(keys, _p(vals), len(params), name.encode(), path.encode(), ptrs, _p(shapes), len(ins), _p(out), _p(oshape)))
    n = int(np.prod(oshape))
    b, h, s, d = [int(v) for v in oshape]
    return out[:n].reshape(b, s, h, d)  # BSHD memory order
