// vision.cu -- the remaining small ops of the multimodal graphs: GELU / QuickGELU
// through the reference's fp16 lookup table, the eager-attention softmax, scalar division, the ViT's 2-D rotary
// embedding, a general strided 4-D copy (transpose / cat / split / head replication) and the load-time
// de-interleave of Q4_0_4_4 weights.  All HBM-bound single passes.
#include "common.cuh"

// ------------------------------------------------------------------ GELU
// mllm_vec_gelu_f32 / mllm_vec_gelu_quick_f32 (ggml/Quantize.hpp:101-133): y = fp32(table[fp16(x)]).  The 128 KB table is
// built by the host exactly as init_table_gelu_f16 does and stays L2-resident.
// fp16 -> fp32 as F16C does it: NaNs keep sign and payload (quieted); cvt.f32.f16 would canonicalise them
static __device__ __forceinline__ float table_f32(const __half *table, float x) {
    const uint32_t u = __half_as_ushort(table[__half_as_ushort(__float2half_rn(x))]);
    if ((u & 0x7C00u) == 0x7C00u && (u & 0x3FFu)) return __uint_as_float(((u & 0x8000u) << 16) | 0x7FC00000u | ((u & 0x3FFu) << 13));
    return __half2float(__ushort_as_half((unsigned short)u));
}
__global__ void k_gelu_table(const float *__restrict__ x, float *__restrict__ y, size_t n, const __half *__restrict__ table) {
    const size_t i = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) * 4;
    if (i >= n) return;
    if (i + 4 <= n) {
        const float4 v = *(const float4 *)(x + i);
        float4 o;
        o.x = table_f32(table, v.x);
        o.y = table_f32(table, v.y);
        o.z = table_f32(table, v.z);
        o.w = table_f32(table, v.w);
        *(float4 *)(y + i) = o;
    } else {
        for (size_t j = i; j < n; j++) y[j] = table_f32(table, x[j]);
    }
}
MLLMCU_API int mllmcu_gelu_table(const float *x, float *y, size_t n, const void *table_f16, void *stream) {
    if (n == 0) return 0;
    if (((uintptr_t)x | (uintptr_t)y) & 15) return -1;
    const size_t thr = (n + 3) / 4;
    k_gelu_table<<<(unsigned)((thr + 255) / 256), 256, 0, (cudaStream_t)stream>>>(x, y, n, (const __half *)table_f16);
    CU_LAUNCH_CHECK();
    return 0;
}

// --------------------------------------------------------------- softmax
// CPUSoftMax::execute (op/CPUSoftMax.cpp:29-66) + mllm_vec_soft_max_f32 (compute/ActivationFunction.cpp:29-80, AVX2 body):
// one warp per row.  Groups of 8 use the polynomial exp and are summed in the AVX lane order
// ((v0+v4)+(v2+v6)) + ((v1+v5)+(v3+v7)); group sums are added serially in order (they are the reference's `sum +=`
// chain, kept so the result is bit-identical); the n%8 tail uses a correctly rounded exp
// (libm's expf is within 0.502 ulp, so ~0.1 % of tail elements differ by 1 ulp; rows with n % 8 == 0 are bit-identical).
__global__ void __launch_bounds__(128) k_softmax(const float *__restrict__ x, float *__restrict__ y, long long rows, int S, int n,
                                                 int causal, int old_dim) {
    extern __shared__ float s_gsum[]; // [4 warps][n/8]
    const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
    const long long row = (long long)blockIdx.x * 4 + wib;
    if (row >= rows) return;
    const int s = (int)(row % S);
    int vis = (causal && S > 1) ? s + 1 + old_dim : n;
    if (vis > n) vis = n;
    const float *xr = x + (size_t)row * n;
    float *yr = y + (size_t)row * n;
    float mx = -INFINITY;
    for (int j = lane; j < vis; j += 32) mx = fmaxf(mx, xr[j]);
    mx = warp_max(mx);
    const int ng = vis / 8;
    float *gs = s_gsum + (size_t)wib * (n / 8 + 1);
    // 4 groups of 8 lanes per pass
    for (int g0 = 0; g0 < ng; g0 += 4) {
        const int g = g0 + (lane >> 3);
        float v = 0.f;
        if (g < ng) { v = v_expf(__fsub_rn(xr[g * 8 + (lane & 7)], mx)); yr[g * 8 + (lane & 7)] = v; }
        const float a = __fadd_rn(__shfl_xor_sync(0xffffffffu, v, 4), v);  // lanes 0..3: v[j+4] + v[j]
        const float b = __fadd_rn(a, __shfl_xor_sync(0xffffffffu, a, 2));  // lanes 0,1: a0+a2, a1+a3
        const float c = __fadd_rn(b, __shfl_xor_sync(0xffffffffu, b, 1));  // lane 0: b0 + b1
        if ((lane & 7) == 0 && g < ng) gs[g] = c;
    }
    __syncwarp();
    float inv = 0.f;
    if (lane == 0) {
        float sum = 0.f;
        for (int g = 0; g < ng; g++) sum = __fadd_rn(sum, gs[g]);
        for (int j = ng * 8; j < vis; j++) {
            const float val = (float)exp((double)__fsub_rn(xr[j], mx));
            sum = __fadd_rn(sum, val);
            yr[j] = val;
        }
        inv = (float)(1.0 / (double)sum);
    }
    inv = __shfl_sync(0xffffffffu, inv, 0);
    __syncwarp();
    for (int j = lane; j < n; j += 32) yr[j] = j < vis ? __fmul_rn(yr[j], inv) : 0.f;
}
MLLMCU_API int mllmcu_softmax(const float *x, float *y, int outer, int S, int n, int causal, int old_dim, void *stream) {
    if (outer <= 0 || S <= 0 || n <= 0 || old_dim < 0) return -1;
    const long long rows = (long long)outer * S;
    const size_t sh = 4 * (size_t)(n / 8 + 1) * sizeof(float);
    if (sh > 48 * 1024) return -2;
    k_softmax<<<(unsigned)((rows + 3) / 4), 128, sh, (cudaStream_t)stream>>>(x, y, rows, S, n, causal, old_dim);
    CU_LAUNCH_CHECK();
    return 0;
}

// ------------------------------------------------------- scalar division
// q / sqrt(D) of the eager attention spelling (Tensor::operator/(float) -> CPUdivFunction, op/CPUBinaryFunc.hpp)
__global__ void k_div_scalar(const float *__restrict__ x, float s, float *__restrict__ y, size_t n) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) y[i] = __fdiv_rn(x[i], s);
}
MLLMCU_API int mllmcu_div_scalar(const float *x, float s, float *y, size_t n, void *stream) {
    if (n == 0) return 0;
    k_div_scalar<<<(unsigned)((n + 255) / 256), 256, 0, (cudaStream_t)stream>>>(x, s, y, n);
    CU_LAUNCH_CHECK();
    return 0;
}

// ----------------------------------------------------------- vision RoPE
// CPUVisionRoPEFuncFunction::rope_hf (op/CPUVisionRoPEFunc.hpp:27-90): rotate-half with sin / cos of angles[s][d], d < D/2.
// table == 0: sin/cos evaluated here (sinf/cosf, <= 2 ulp from libm); table == 1: `angles` and `angles2` already hold
// host-built sin and cos ([S][D/2] each), which makes the result bit-identical to the reference.
__global__ void k_rope_vision(const float *__restrict__ x, const float *__restrict__ a0, const float *__restrict__ a1, void *out,
                              long long total, int S, int H, int D, int table, int out_f16) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= total) return;
    const int half = D / 2;
    const int d = (int)(i % half);
    const int h = (int)((i / half) % H);
    const long long bs = i / ((long long)half * H); // b*S + s
    const int s = (int)(bs % S);
    const size_t base = ((size_t)bs * H + h) * D + d;
    float sn, cs;
    if (table) { sn = a0[(size_t)s * half + d]; cs = a1[(size_t)s * half + d]; }
    else { const float a = a0[(size_t)s * half + d]; sn = sinf(a); cs = cosf(a); }
    const float v0 = x[base], v1 = x[base + half];
    const float o0 = rope_lo(v0, v1, sn, cs);
    const float o1 = rope_hi(v0, v1, sn, cs);
    if (out_f16) { ((__half *)out)[base] = __float2half_rn(o0); ((__half *)out)[base + half] = __float2half_rn(o1); }
    else { ((float *)out)[base] = o0; ((float *)out)[base + half] = o1; }
}
MLLMCU_API int mllmcu_rope_vision(const float *x, const float *angles_or_sin, const float *cos_or_null, void *out, int B, int S, int H,
                                  int D, int out_f16, void *stream) {
    if (B <= 0 || S <= 0 || H <= 0 || D <= 0 || D % 2) return -1;
    const long long total = (long long)B * S * H * (D / 2);
    k_rope_vision<<<(unsigned)((total + 255) / 256), 256, 0, (cudaStream_t)stream>>>(x, angles_or_sin, cos_or_null, out, total, S, H, D,
                                                                                      cos_or_null != nullptr, out_f16);
    CU_LAUNCH_CHECK();
    return 0;
}

// ----------------------------------------------------------- strided copy
// One kernel behind the reference's physical data movers: CPUtransposeFunction (op/CPUTransposeFunc.hpp:127-149),
// CPUcatFunction, CPUsplitFunction, the n_rep head replication of CPUKVCache (op/CPUKVCache.cpp:137-247), clip copies.
// dst[i0*ds0 + i1*ds1 + i2*ds2 + i3*ds3] = src[i0*ss0 + ...] over an n0 x n1 x n2 x n3 index space (strides in elements).
template <typename T>
__global__ void k_copy_strided(const T *__restrict__ src, T *__restrict__ dst, int n1, int n2, int n3, long long total, long long ss0,
                               long long ss1, long long ss2, long long ss3, long long ds0, long long ds1, long long ds2, long long ds3) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= total) return;
    const int i3 = (int)(i % n3);
    const int i2 = (int)((i / n3) % n2);
    const int i1 = (int)((i / ((long long)n3 * n2)) % n1);
    const long long i0 = i / ((long long)n3 * n2 * n1);
    dst[i0 * ds0 + i1 * ds1 + i2 * ds2 + i3 * ds3] = src[i0 * ss0 + i1 * ss1 + i2 * ss2 + i3 * ss3];
}
MLLMCU_API int mllmcu_copy_strided(const void *src, void *dst, int elem_bytes, const int *n, const long long *ss, const long long *ds,
                                   void *stream) {
    if (!n || !ss || !ds || n[0] <= 0 || n[1] <= 0 || n[2] <= 0 || n[3] <= 0) return -1;
    const long long total = (long long)n[0] * n[1] * n[2] * n[3];
    const unsigned grid = (unsigned)((total + 255) / 256);
    cudaStream_t st = (cudaStream_t)stream;
    if (elem_bytes == 4)
        k_copy_strided<uint32_t><<<grid, 256, 0, st>>>((const uint32_t *)src, (uint32_t *)dst, n[1], n[2], n[3], total, ss[0], ss[1], ss[2], ss[3], ds[0], ds[1], ds[2], ds[3]);
    else if (elem_bytes == 2)
        k_copy_strided<uint16_t><<<grid, 256, 0, st>>>((const uint16_t *)src, (uint16_t *)dst, n[1], n[2], n[3], total, ss[0], ss[1], ss[2], ss[3], ds[0], ds[1], ds[2], ds[3]);
    else return -2;
    CU_LAUNCH_CHECK();
    return 0;
}

// ---------------------------------------------------- Q4_0_4_4 de-interleave
// The shipped *-q4_0_4_4.mllm files hold block_q4_0x4 groups (make_block_q4_0x4, ggml/GemmPack.cpp:45-60, interleave 4, nibbles
// XOR 0x88; rows 4g..4g+3 of block x at ((g*nb)+x)*72).  At load they are rewritten once into plain block_q4_0 rows, the layout
// every Linear kernel here streams (a load-time repack, not a kernel family).  One thread per output qs byte.
__global__ void k_q4_0_4x4_unpack(const uint8_t *__restrict__ packed, uint8_t *__restrict__ rows, long long total, int nb) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= total) return;
    const int j = (int)(i % 18);           // byte inside the output block (0,1 = d)
    const long long blk = i / 18;          // output block index: row*nb + x
    const int x = (int)(blk % nb);
    const long long row = blk / nb;
    const int r = (int)(row & 3);
    const uint8_t *p = packed + ((row >> 2) * nb + x) * 72;
    uint8_t v;
    if (j < 2) v = p[2 * r + j];
    else { const int q = j - 2; v = p[8 + (q >> 2) * 16 + r * 4 + (q & 3)] ^ 0x88; }
    rows[i] = v;
}
MLLMCU_API int mllmcu_q4_0_4x4_to_q4_0(const void *packed, void *rows, int nrow, int k, void *stream) {
    if (nrow <= 0 || nrow % 4 || k <= 0 || k % QK || packed == rows) return -1;
    const long long total = (long long)nrow * (k / QK) * 18;
    k_q4_0_4x4_unpack<<<(unsigned)((total + 255) / 256), 256, 0, (cudaStream_t)stream>>>((const uint8_t *)packed, (uint8_t *)rows, total, k / QK);
    CU_LAUNCH_CHECK();
    return 0;
}
