// elementwise.cu -- the HBM-bound small ops of the decoder block: RMSNorm,
// LayerNorm, RoPE, SiLU / SiLU*mul, add / mul, embedding gather-dequant,
// fp32->fp16 rows, argmax.  Each is a single pass over its data with coalesced
// (vectorised where alignment allows) accesses; reference citations per kernel.
#include "common.cuh"

// ------------------------------------------------------------------ RMSNorm
// CPURMSNorm::execute (op/CPURMSNorm.cpp:31-95): sum of squares accumulated in
// double, mean/rms in float, then two separately rounded multiplies.
__global__ void __launch_bounds__(256) k_rmsnorm(const float *__restrict__ x, const float *__restrict__ w, float *__restrict__ y,
                                                 int dim, float eps, int add_unit_offset) {
    const int r = blockIdx.x;
    const float *xr = x + (size_t)r * dim;
    float *yr = y + (size_t)r * dim;
    const bool vec = (dim & 3) == 0 && ((((uintptr_t)xr) | ((uintptr_t)yr) | ((uintptr_t)w)) & 15) == 0; // 16-byte loads and stores
    double ss = 0.0;
    if (vec) {
        const float4 *x4 = (const float4 *)xr;
        double s0 = 0.0, s1 = 0.0;
        for (int d = threadIdx.x; d < (dim >> 2); d += blockDim.x) {
            const float4 v = x4[d];
            s0 = fma((double)v.x, (double)v.x, s0); s1 = fma((double)v.y, (double)v.y, s1);
            s0 = fma((double)v.z, (double)v.z, s0); s1 = fma((double)v.w, (double)v.w, s1);
        }
        ss = s0 + s1;
    } else {
        for (int d = threadIdx.x; d < dim; d += blockDim.x) {
            const float v = xr[d];
            ss += (double)v * (double)v;
        }
    }
    __shared__ double sred[8];
    __shared__ float s_rms;
    ss = warp_sum_d(ss);
    if ((threadIdx.x & 31) == 0) sred[threadIdx.x >> 5] = ss;
    __syncthreads();
    if (threadIdx.x == 0) {
        double t = 0.0;
        for (int i = 0; i < (int)(blockDim.x >> 5); i++) t += sred[i];
        const float mean = (float)(t / (double)dim);
        s_rms = __fdiv_rn(1.0f, __fsqrt_rn(__fadd_rn(mean, eps)));
    }
    __syncthreads();
    const float rms = s_rms;
    if (vec) {
        const float4 *x4 = (const float4 *)xr, *w4 = (const float4 *)w;
        float4 *y4 = (float4 *)yr;
        for (int d = threadIdx.x; d < (dim >> 2); d += blockDim.x) {
            const float4 v = x4[d];
            float4 ww = w4[d];
            if (add_unit_offset) { ww.x = __fadd_rn(1.0f, ww.x); ww.y = __fadd_rn(1.0f, ww.y); ww.z = __fadd_rn(1.0f, ww.z); ww.w = __fadd_rn(1.0f, ww.w); }
            y4[d] = make_float4(__fmul_rn(__fmul_rn(v.x, rms), ww.x), __fmul_rn(__fmul_rn(v.y, rms), ww.y), __fmul_rn(__fmul_rn(v.z, rms), ww.z),
                                __fmul_rn(__fmul_rn(v.w, rms), ww.w));
        }
        return;
    }
    for (int d = threadIdx.x; d < dim; d += blockDim.x) {
        const float t = __fmul_rn(xr[d], rms);
        const float ww = add_unit_offset ? __fadd_rn(1.0f, w[d]) : w[d];
        yr[d] = __fmul_rn(t, ww);
    }
}

MLLMCU_API int mllmcu_rmsnorm(const float *x, const float *w, float *y, int rows, int dim, float eps, int add_unit_offset, void *stream) {
    if (rows <= 0 || dim <= 0) return -1;
    k_rmsnorm<<<rows, 256, 0, (cudaStream_t)stream>>>(x, w, y, dim, eps, add_unit_offset);
    CU_LAUNCH_CHECK();
    return 0;
}

// ---------------------------------------------------------------- LayerNorm
// CPULayerNorm::execute (op/CPULayerNorm.cpp:49-89): two-pass mean / sum of squared deviations.
__global__ void __launch_bounds__(256) k_layernorm(const float *__restrict__ x, const float *__restrict__ w, const float *__restrict__ b,
                                                   float *__restrict__ y, int dim, float eps) {
    const int r = blockIdx.x;
    const float *xr = x + (size_t)r * dim;
    float *yr = y + (size_t)r * dim;
    __shared__ float sred[8];
    __shared__ float s_val;
    float s = 0.f;
    for (int d = threadIdx.x; d < dim; d += blockDim.x) s += xr[d];
    s = warp_sum(s);
    if ((threadIdx.x & 31) == 0) sred[threadIdx.x >> 5] = s;
    __syncthreads();
    if (threadIdx.x == 0) {
        float t = 0.f;
        for (int i = 0; i < (int)(blockDim.x >> 5); i++) t += sred[i];
        s_val = t / dim;
    }
    __syncthreads();
    const float mean = s_val;
    float q = 0.f;
    for (int d = threadIdx.x; d < dim; d += blockDim.x) { const float t = xr[d] - mean; q += t * t; }
    q = warp_sum(q);
    __syncthreads();
    if ((threadIdx.x & 31) == 0) sred[threadIdx.x >> 5] = q;
    __syncthreads();
    if (threadIdx.x == 0) {
        float t = 0.f;
        for (int i = 0; i < (int)(blockDim.x >> 5); i++) t += sred[i];
        s_val = __fsqrt_rn(__fadd_rn(__fdiv_rn(t, (float)dim), eps));
    }
    __syncthreads();
    const float rms = s_val;
    // the reference's element formula, in its order: w * (x - mean) / rms + b -- a multiply, a DIVISION, an add (:77-81).
    // Only the two sums differ from it (tree order here, serial there): results within a few ulp, tested against the compiled op.
    for (int d = threadIdx.x; d < dim; d += blockDim.x) {
        float v = __fsub_rn(xr[d], mean);
        if (w) v = __fmul_rn(w[d], v);
        v = __fdiv_rn(v, rms);
        if (b) v = __fadd_rn(v, b[d]);
        yr[d] = v;
    }
}

MLLMCU_API int mllmcu_layernorm(const float *x, const float *w, const float *b, float *y, int rows, int dim, float eps, void *stream) {
    if (rows <= 0 || dim <= 0) return -1;
    k_layernorm<<<rows, 256, 0, (cudaStream_t)stream>>>(x, w, b, y, dim, eps);
    CU_LAUNCH_CHECK();
    return 0;
}

// --------------------------------------------------------------------- RoPE
// CPURoPE::rope_llama / rope_hf (op/CPURoPE.cpp:200-372): one thread per rotated
// pair (or per pass-through element beyond rot_dim).
__global__ void k_rope(int type, const float *__restrict__ x, void *out, int B, int S, int H, int D, int rot, int pos0,
                       const float *__restrict__ sin_t, const float *__restrict__ cos_t, int tdim, int out_f16,
                       long long obs, long long oss) {
    const int half_d = D / 2; // threads per (b,s,h): D/2 (pairs) -- pass-through dims handled pairwise too
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    const long long total = (long long)B * S * H * half_d;
    if (i >= total) return;
    const int p = (int)(i % half_d);
    const int h = (int)((i / half_d) % H);
    const int s = (int)((i / ((long long)half_d * H)) % S);
    const int b = (int)(i / ((long long)half_d * H * S));
    const float *xi = x + (((size_t)b * S + s) * H + h) * D;
    const size_t ob = (size_t)b * obs + (size_t)s * oss + (size_t)h * D;
    int d0, d1;
    float v0, v1;
    const int hr = rot / 2;
    if (p < hr) {
        if (type == 2) { d0 = 2 * p; d1 = 2 * p + 1; }
        else { d0 = p; d1 = p + hr; }
        const float a = xi[d0], bb = xi[d1];
        const float sn = sin_t[(size_t)(s + pos0) * tdim + d0], cs = cos_t[(size_t)(s + pos0) * tdim + d0];
        // a*cos - b*sin, a*sin + b*cos; products rounded separately (host build uses no FMA here: tolerance noted in tests)
        v0 = rope_lo(a, bb, sn, cs);
        v1 = rope_hi(a, bb, sn, cs);
    } else { // beyond the rotated part: copy through
        d0 = rot + 2 * (p - hr); d1 = d0 + 1;
        v0 = xi[d0]; v1 = xi[d1];
    }
    if (out_f16) {
        ((__half *)out)[ob + d0] = __float2half_rn(v0);
        ((__half *)out)[ob + d1] = __float2half_rn(v1);
    } else {
        ((float *)out)[ob + d0] = v0;
        ((float *)out)[ob + d1] = v1;
    }
}

// LLaMA-type rotation (adjacent pairs) over the whole head: two pairs per thread, 16-byte loads of x and of the table rows.  Same
// operations per element as k_rope (products rounded separately), so the results are bit-identical.
__global__ void k_rope_llama4(const float *__restrict__ x, void *out, int B, int S, int H, int D, int pos0, const float *__restrict__ sin_t,
                              const float *__restrict__ cos_t, int tdim, int out_f16, long long obs, long long oss) {
    const int q4 = D / 4;
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (long long)B * S * H * q4) return;
    const int p = (int)(i % q4);
    const int h = (int)((i / q4) % H);
    const int s = (int)((i / ((long long)q4 * H)) % S);
    const int b = (int)(i / ((long long)q4 * H * S));
    const float4 v = *(const float4 *)(x + (((size_t)b * S + s) * H + h) * D + 4 * p);
    const float4 sn = *(const float4 *)(sin_t + (size_t)(s + pos0) * tdim + 4 * p), cs = *(const float4 *)(cos_t + (size_t)(s + pos0) * tdim + 4 * p);
    const float o0 = rope_lo(v.x, v.y, sn.x, cs.x), o1 = rope_hi(v.x, v.y, sn.x, cs.x);
    const float o2 = rope_lo(v.z, v.w, sn.z, cs.z), o3 = rope_hi(v.z, v.w, sn.z, cs.z);
    const size_t at = (size_t)b * obs + (size_t)s * oss + (size_t)h * D + 4 * p;
    if (out_f16) {
        const __half2 a = __halves2half2(__float2half_rn(o0), __float2half_rn(o1)), c = __halves2half2(__float2half_rn(o2), __float2half_rn(o3));
        *(uint2 *)((__half *)out + at) = make_uint2(*(const uint32_t *)&a, *(const uint32_t *)&c);
    } else {
        *(float4 *)((float *)out + at) = make_float4(o0, o1, o2, o3);
    }
}

MLLMCU_API int mllmcu_rope(int rope_type, const float *x, void *out, int B, int S, int H, int D, int rot_dim, int pos0,
                           const float *sin_t, const float *cos_t, int tdim, int out_f16, long long obs, long long oss, void *stream) {
    if ((rope_type != 2 && rope_type != 4) || D % 2 || rot_dim % 2 || rot_dim > D || rot_dim > tdim) return -1;
    const long long total = (long long)B * S * H * (D / 2);
    if (total <= 0) return -1;
    const size_t esz = out_f16 ? 2 : 4;
    if (rope_type == 2 && rot_dim == D && D % 4 == 0 && tdim % 4 == 0 && obs % 4 == 0 && oss % 4 == 0 && total >= 4096 &&
        !((uintptr_t)x & 15) && !((uintptr_t)sin_t & 15) && !((uintptr_t)cos_t & 15) && !((uintptr_t)out & (4 * esz - 1))) {
        const long long t4 = total / 2;
        k_rope_llama4<<<(unsigned)((t4 + 255) / 256), 256, 0, (cudaStream_t)stream>>>(x, out, B, S, H, D, pos0, sin_t, cos_t, tdim, out_f16, obs, oss);
        CU_LAUNCH_CHECK();
        return 0;
    }
    k_rope<<<(unsigned)((total + 255) / 256), 256, 0, (cudaStream_t)stream>>>(rope_type, x, out, B, S, H, D, rot_dim, pos0, sin_t, cos_t, tdim, out_f16, obs, oss);
    CU_LAUNCH_CHECK();
    return 0;
}

MLLMCU_API int mllmcu_rope_mrope(const float *x, void *out, int B, int S, int H, int D, int rot_dim, const float *sin_t, const float *cos_t,
                                 int tdim, int out_f16, long long obs, long long oss, void *stream) {
    return mllmcu_rope(4, x, out, B, S, H, D, rot_dim, 0, sin_t, cos_t, tdim, out_f16, obs, oss, stream);
}

// --------------------------------------------------------------------- SiLU
// mllm_v_expf / mllm_v_silu (compute/ActivationFunction.hpp:96-141) restated per
// element with the same fused / unfused operations, so results are bit-identical
// to the AVX2 (and AVX-512) vector bodies.
static __device__ __forceinline__ float silu_f(float x) {
    return __fdiv_rn(x, __fadd_rn(1.0f, v_expf(__fsub_rn(0.0f, x))));
}

template <int MODE> // 0 silu, 1 silu(a)*b, 2 a+b, 3 a*b
__global__ void k_ew(const float *__restrict__ a, const float *__restrict__ b, float *__restrict__ y, size_t n) {
    const size_t i = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) * 4;
    if (i >= n) return;
    const bool vec = (i + 4 <= n);
    float av[4], bv[4] = {0, 0, 0, 0}, o[4];
    if (vec) {
        const float4 t = *(const float4 *)(a + i);
        av[0] = t.x; av[1] = t.y; av[2] = t.z; av[3] = t.w;
        if (MODE != 0) { const float4 u = *(const float4 *)(b + i); bv[0] = u.x; bv[1] = u.y; bv[2] = u.z; bv[3] = u.w; }
    } else {
        for (int j = 0; j < 4; j++) { av[j] = (i + j < n) ? a[i + j] : 0.f; if (MODE != 0) bv[j] = (i + j < n) ? b[i + j] : 0.f; }
    }
#pragma unroll
    for (int j = 0; j < 4; j++) {
        if (MODE == 0) o[j] = silu_f(av[j]);
        else if (MODE == 1) o[j] = __fmul_rn(silu_f(av[j]), bv[j]);
        else if (MODE == 2) o[j] = __fadd_rn(av[j], bv[j]);
        else o[j] = __fmul_rn(av[j], bv[j]);
    }
    if (vec) *(float4 *)(y + i) = make_float4(o[0], o[1], o[2], o[3]);
    else for (int j = 0; j < 4 && i + j < n; j++) y[i + j] = o[j];
}

template <int MODE>
static int launch_ew(const float *a, const float *b, float *y, size_t n, void *stream) {
    if (n == 0) return 0;
    if (((uintptr_t)a | (uintptr_t)y | (uintptr_t)(MODE ? b : a)) & 15) return -1;
    const size_t thr = (n + 3) / 4;
    k_ew<MODE><<<(unsigned)((thr + 255) / 256), 256, 0, (cudaStream_t)stream>>>(a, b, y, n);
    CU_LAUNCH_CHECK();
    return 0;
}
MLLMCU_API int mllmcu_silu(const float *x, float *y, size_t n, void *stream) { return launch_ew<0>(x, nullptr, y, n, stream); }
MLLMCU_API int mllmcu_silu_mul(const float *g, const float *u, float *y, size_t n, void *stream) { return launch_ew<1>(g, u, y, n, stream); }
MLLMCU_API int mllmcu_add(const float *a, const float *b, float *y, size_t n, void *stream) { return launch_ew<2>(a, b, y, n, stream); }
MLLMCU_API int mllmcu_mul(const float *a, const float *b, float *y, size_t n, void *stream) { return launch_ew<3>(a, b, y, n, stream); }

// ---------------------------------------------------------------- Embedding
// CPUEmbedding::execute Q4_0 (op/CPUEmbedding.cpp:66-81) + dequantize_row_q4_0
// (ggml/QuantizeQ4.cpp:74-93).  One thread per (token, block, byte j): writes
// elements j and j+16 of the block.
__global__ void k_embedding_q4_0(const uint8_t *__restrict__ table, const float *__restrict__ ids, float *__restrict__ out, int n_ids, int hidden) {
    const int nb = hidden / QK;
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (long long)n_ids * nb * 16) return;
    const int j = (int)(i & 15);
    const int b = (int)((i >> 4) % nb);
    const int t = (int)((i >> 4) / nb);
    const float idf = ids[t];
    if (!(idf >= 0)) return;
    const uint8_t *blk = table + ((size_t)(int)idf * nb + b) * 18;
    const float d = __half2float(*(const __half *)blk);
    const uint8_t q = blk[2 + j];
    float *o = out + (size_t)t * hidden + (size_t)b * QK;
    o[j] = __fmul_rn((float)((int)(q & 0x0F) - 8), d);
    o[j + 16] = __fmul_rn((float)((int)(q >> 4) - 8), d);
}
MLLMCU_API int mllmcu_embedding_q4_0(const void *table, const float *ids, float *out, int n_ids, int hidden, void *stream) {
    if (n_ids <= 0 || hidden % QK) return -1;
    const long long tot = (long long)n_ids * (hidden / QK) * 16;
    k_embedding_q4_0<<<(unsigned)((tot + 255) / 256), 256, 0, (cudaStream_t)stream>>>((const uint8_t *)table, ids, out, n_ids, hidden);
    CU_LAUNCH_CHECK();
    return 0;
}

// CPUEmbedding::execute, F32 table (op/CPUEmbedding.cpp:44-63): a row copy per token; negative ids leave the row untouched
__global__ void k_embedding_f32(const float *__restrict__ table, const float *__restrict__ ids, float *__restrict__ out, int n_ids, int hidden) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (long long)n_ids * hidden) return;
    const int t = (int)(i / hidden), d = (int)(i % hidden);
    const float idf = ids[t];
    if (!(idf >= 0)) return;
    out[i] = table[(size_t)(int)idf * hidden + d];
}
MLLMCU_API int mllmcu_embedding_f32(const void *table, const float *ids, float *out, int n_ids, int hidden, void *stream) {
    if (n_ids <= 0 || hidden <= 0) return -1;
    const long long tot = (long long)n_ids * hidden;
    k_embedding_f32<<<(unsigned)((tot + 255) / 256), 256, 0, (cudaStream_t)stream>>>((const float *)table, ids, out, n_ids, hidden);
    CU_LAUNCH_CHECK();
    return 0;
}

// ------------------------------------------------------------- fp32 -> fp16
__global__ void k_f32_to_f16_rows(const float *__restrict__ src, __half *__restrict__ dst, int rows, int cols, long long lds, long long ldd) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (long long)rows * cols) return;
    const int c = (int)(i % cols);
    const long long r = i / cols;
    dst[r * ldd + c] = __float2half_rn(src[r * lds + c]);
}
MLLMCU_API int mllmcu_f32_to_f16_rows(const float *src, void *dst, int rows, int cols, long long lds, long long ldd, void *stream) {
    if (rows <= 0 || cols <= 0) return -1;
    const long long tot = (long long)rows * cols;
    k_f32_to_f16_rows<<<(unsigned)((tot + 255) / 256), 256, 0, (cudaStream_t)stream>>>(src, (__half *)dst, rows, cols, lds, ldd);
    CU_LAUNCH_CHECK();
    return 0;
}

// fp16 -> fp32 (Backend::convert_fp_data, Tensor::fp32(): exact widening)
__global__ void k_f16_to_f32_rows(const __half *__restrict__ src, float *__restrict__ dst, int rows, int cols, long long lds, long long ldd) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (long long)rows * cols) return;
    const int c = (int)(i % cols);
    const long long r = i / cols;
    dst[r * ldd + c] = __half2float(src[r * lds + c]);
}
MLLMCU_API int mllmcu_f16_to_f32_rows(const void *src, float *dst, int rows, int cols, long long lds, long long ldd, void *stream) {
    if (rows <= 0 || cols <= 0) return -1;
    const long long tot = (long long)rows * cols;
    k_f16_to_f32_rows<<<(unsigned)((tot + 255) / 256), 256, 0, (cudaStream_t)stream>>>((const __half *)src, dst, rows, cols, lds, ldd);
    CU_LAUNCH_CHECK();
    return 0;
}

// ------------------------------------------------------------------- argmax
// LlmTextGenerator greedy (Generate.cpp:46-53): first maximum wins.
__global__ void __launch_bounds__(1024) k_argmax(const float *__restrict__ x, int n, int *__restrict__ out, float *__restrict__ out_f) {
    const float *xr = x + (size_t)blockIdx.x * n;
    float best = -INFINITY;
    int bi = 0x7fffffff;
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
        const float v = xr[i];
        if (v > best || (v == best && i < bi)) { best = v; bi = i; }
    }
    __shared__ float sv[32];
    __shared__ int si[32];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const float ov = __shfl_xor_sync(0xffffffffu, best, o);
        const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
        if (ov > best || (ov == best && oi < bi)) { best = ov; bi = oi; }
    }
    if ((threadIdx.x & 31) == 0) { sv[threadIdx.x >> 5] = best; si[threadIdx.x >> 5] = bi; }
    __syncthreads();
    if (threadIdx.x < 32) {
        const int nw = blockDim.x >> 5;
        best = threadIdx.x < nw ? sv[threadIdx.x] : -INFINITY;
        bi = threadIdx.x < nw ? si[threadIdx.x] : 0x7fffffff;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            const float ov = __shfl_xor_sync(0xffffffffu, best, o);
            const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
            if (ov > best || (ov == best && oi < bi)) { best = ov; bi = oi; }
        }
        if (threadIdx.x == 0) {
            const int r = (bi == 0x7fffffff) ? 0 : bi;
            if (out) out[blockIdx.x] = r;
            if (out_f) out_f[blockIdx.x] = (float)r;
        }
    }
}
MLLMCU_API int mllmcu_argmax(const float *logits, int rows, int n, int *out, void *stream) {
    if (rows <= 0 || n <= 0) return -1;
    k_argmax<<<rows, 1024, 0, (cudaStream_t)stream>>>(logits, n, out, nullptr);
    CU_LAUNCH_CHECK();
    return 0;
}
MLLMCU_API int mllmcu_argmax_ids(const float *logits, int rows, int n, float *ids_out, void *stream) {
    if (rows <= 0 || n <= 0) return -1;
    k_argmax<<<rows, 1024, 0, (cudaStream_t)stream>>>(logits, n, nullptr, ids_out);
    CU_LAUNCH_CHECK();
    return 0;
}
