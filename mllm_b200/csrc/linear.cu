// linear.cu -- block-quantised Linear (Q4_0 / Q8_0 weights x Q8_0 activations).
// Reference: CPULinear::execute -> mat_mul (op/CPULinear.cpp:97-233,
// compute/Matmul.cpp:77-120 activation quantisation, :219-276 vec_dot loop),
// vec_dot_q4_0_q8_0 (ggml/VecDotQ4.cpp:514-544), vec_dot_q8_0_q8_0 (ggml/VecDotQ8.cpp:127-149).
//
// Weights stay in the .mllm file's native block layout in HBM: row n is K/32
// consecutive 18-byte block_q4_0 (or 34-byte block_q8_0).  Activations are
// quantised once per call into a split layout (int8 qs rows + fp32 scales).
// The integer block dot products are exact; the fp32 scale-accumulate order
// differs from the CPU's 8-lane AVX pattern, so fp32 outputs agree to
// tolerance (tests state it), as north_star allows.
//
// Two paths:
//   M <= GEMV_MAX_M : k_gemv_*  -- HBM-bound weight streaming (decode)
//   otherwise       : k_gemm_*  -- tiled integer GEMM (prefill / large batch)
#include <cstdio>
#include "common.cuh"
#include "quantize.cuh"
#include "mllmcu.h"

#define GEMV_MAX_M 8

static __device__ __forceinline__ uint32_t ld_u32_a2(const uint8_t *p) {
    // 4 bytes from a 2-byte-aligned address
    const uint16_t *q = (const uint16_t *)p;
    return (uint32_t)q[0] | ((uint32_t)q[1] << 16);
}

// integer dot of one Q4_0 block (16 qs bytes at 2-byte alignment) with 32 int8
// activations a[0..7] (words): sum (q-8)*a, exact.
static __device__ __forceinline__ int q4_0_block_idot(const uint8_t *qs, const int *a) {
    int s = 0, sa = 0;
#pragma unroll
    for (int j = 0; j < 4; j++) {
        const uint32_t w = ld_u32_a2(qs + 4 * j);
        const int lo = (int)(w & 0x0F0F0F0Fu), hi = (int)((w >> 4) & 0x0F0F0F0Fu);
        s = __dp4a(lo, a[j], s);
        s = __dp4a(hi, a[4 + j], s);
        sa = __dp4a(0x01010101, a[j], sa);
        sa = __dp4a(0x01010101, a[4 + j], sa);
    }
    return s - 8 * sa;
}

static __device__ __forceinline__ int q8_0_block_idot(const int8_t *qs, const int *a) {
    int s = 0;
#pragma unroll
    for (int j = 0; j < 8; j++) s = __dp4a((int)ld_u32_a2((const uint8_t *)qs + 4 * j), a[j], s);
    return s;
}

static __device__ __forceinline__ void store_out(void *y, size_t idx, float v, int out_f16) {
    if (out_f16) ((__half *)y)[idx] = __float2half_rn(v);
    else ((float *)y)[idx] = v;
}

// ---------------------------------------------------------------- GEMV v1
// one warp per output row; lane l takes blocks l, l+32, ...
template <int M, int WT>
__global__ void __launch_bounds__(128) k_gemv(const uint8_t *__restrict__ W, const int8_t *__restrict__ xq,
                                              const float *__restrict__ xd, const float *__restrict__ bias, void *y, int N,
                                              int K, int ldy, int out_f16) {
    const int lane = threadIdx.x & 31;
    const int n = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (n >= N) return;
    const int nb = K / QK;
    constexpr int BS = (WT == 2) ? 18 : 34;
    const uint8_t *row = W + (size_t)n * nb * BS;
    float acc[M];
#pragma unroll
    for (int m = 0; m < M; m++) acc[m] = 0.f;
    for (int b = lane; b < nb; b += 32) {
        const uint8_t *blk = row + (size_t)b * BS;
        const float dw = __half2float(*(const __half *)blk);
#pragma unroll
        for (int m = 0; m < M; m++) {
            const int4 *ap = (const int4 *)(xq + (size_t)m * K + (size_t)b * QK);
            const int4 a0 = ap[0], a1 = ap[1];
            const int a[8] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w};
            const int is = (WT == 2) ? q4_0_block_idot(blk + 2, a) : q8_0_block_idot((const int8_t *)blk + 2, a);
            acc[m] = fmaf(dw * xd[(size_t)m * nb + b], (float)is, acc[m]);
        }
    }
#pragma unroll
    for (int m = 0; m < M; m++) {
        float v = warp_sum(acc[m]);
        if (lane == 0) {
            if (bias) v = v + bias[n];
            store_out(y, (size_t)m * ldy + n, v, out_f16);
        }
    }
}

// ---------------------------------------------------------------- GEMM v1
// 64x64 output tile per CTA of 256 threads (4x4 per thread), one 32-wide K
// block per step; weights unpacked to int8 (q-8) in shared memory.
template <int WT>
__global__ void __launch_bounds__(256) k_gemm(const uint8_t *__restrict__ W, const int8_t *__restrict__ xq,
                                              const float *__restrict__ xd, const float *__restrict__ bias, void *y, int M,
                                              int N, int K, int ldy, int out_f16) {
    constexpr int BM = 64, BN = 64;
    constexpr int BS = (WT == 2) ? 18 : 34;
    __shared__ int sA[BM][9]; // 8 words + pad
    __shared__ int sB[BN][9];
    __shared__ float sAd[BM], sBd[BN];
    const int nb = K / QK;
    const int m0 = blockIdx.y * BM, n0 = blockIdx.x * BN;
    const int t = threadIdx.x, tx = t & 15, ty = t >> 4;
    float acc[4][4];
#pragma unroll
    for (int i = 0; i < 4; i++)
#pragma unroll
        for (int j = 0; j < 4; j++) acc[i][j] = 0.f;
    for (int b = 0; b < nb; b++) {
        if (t < 128) { // activation tile: 64 rows x 2 int4
            const int r = t >> 1, h = t & 1;
            int4 v = make_int4(0, 0, 0, 0);
            if (m0 + r < M) v = *(const int4 *)(xq + (size_t)(m0 + r) * K + (size_t)b * QK + h * 16);
            sA[r][h * 4 + 0] = v.x; sA[r][h * 4 + 1] = v.y; sA[r][h * 4 + 2] = v.z; sA[r][h * 4 + 3] = v.w;
            if (h == 0) sAd[r] = (m0 + r < M) ? xd[(size_t)(m0 + r) * nb + b] : 0.f;
        } else if (t < 192) { // weight tile: 64 blocks, one per thread
            const int r = t - 128;
            if (n0 + r < N) {
                const uint8_t *blk = W + ((size_t)(n0 + r) * nb + b) * BS;
                sBd[r] = __half2float(*(const __half *)blk);
                if (WT == 2) {
#pragma unroll
                    for (int j = 0; j < 4; j++) {
                        const uint32_t w = ld_u32_a2(blk + 2 + 4 * j);
                        // (q - 8) per byte: q in 0..15 -> two's complement int8
                        const uint32_t lo = w & 0x0F0F0F0Fu, hi = (w >> 4) & 0x0F0F0F0Fu;
                        sB[r][j] = (int)__vsub4(lo, 0x08080808u);
                        sB[r][4 + j] = (int)__vsub4(hi, 0x08080808u);
                    }
                } else {
#pragma unroll
                    for (int j = 0; j < 8; j++) sB[r][j] = (int)ld_u32_a2(blk + 2 + 4 * j);
                }
            } else {
                sBd[r] = 0.f;
#pragma unroll
                for (int j = 0; j < 8; j++) sB[r][j] = 0;
            }
        }
        __syncthreads();
        int is[4][4];
#pragma unroll
        for (int i = 0; i < 4; i++)
#pragma unroll
            for (int j = 0; j < 4; j++) is[i][j] = 0;
#pragma unroll
        for (int w = 0; w < 8; w++) {
            int a[4], bb[4];
#pragma unroll
            for (int i = 0; i < 4; i++) a[i] = sA[ty * 4 + i][w];
#pragma unroll
            for (int j = 0; j < 4; j++) bb[j] = sB[tx * 4 + j][w];
#pragma unroll
            for (int i = 0; i < 4; i++)
#pragma unroll
                for (int j = 0; j < 4; j++) is[i][j] = __dp4a(a[i], bb[j], is[i][j]);
        }
#pragma unroll
        for (int i = 0; i < 4; i++)
#pragma unroll
            for (int j = 0; j < 4; j++) acc[i][j] = fmaf(sBd[tx * 4 + j] * sAd[ty * 4 + i], (float)is[i][j], acc[i][j]);
        __syncthreads();
    }
#pragma unroll
    for (int i = 0; i < 4; i++) {
        const int m = m0 + ty * 4 + i;
        if (m >= M) continue;
#pragma unroll
        for (int j = 0; j < 4; j++) {
            const int n = n0 + tx * 4 + j;
            if (n >= N) continue;
            float v = acc[i][j];
            if (bias) v = v + bias[n];
            store_out(y, (size_t)m * ldy + n, v, out_f16);
        }
    }
}

// ------------------------------------------------- exact integer block sums
__global__ void k_block_isums_q4_0(const uint8_t *__restrict__ W, const uint8_t *__restrict__ xq, int32_t *__restrict__ out,
                                   int M, int N, int nb) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (long long)M * N * nb) return;
    const int b = (int)(i % nb);
    const int m = (int)((i / nb) % M);
    const int n = (int)(i / ((long long)nb * M));
    const uint8_t *wb = W + ((size_t)n * nb + b) * 18 + 2;
    const uint8_t *xb = xq + ((size_t)m * nb + b) * 34 + 2;
    int a[8];
#pragma unroll
    for (int j = 0; j < 8; j++) a[j] = (int)ld_u32_a2(xb + 4 * j);
    out[i] = q4_0_block_idot(wb, a);
}

MLLMCU_API int mllmcu_block_isums_q4_0_q8_0(const void *W, const void *xq, int32_t *out, int M, int N, int K, void *stream) {
    if (K % QK || M <= 0 || N <= 0) return -1;
    const int nb = K / QK;
    const long long tot = (long long)M * N * nb;
    k_block_isums_q4_0<<<(unsigned)((tot + 255) / 256), 256, 0, (cudaStream_t)stream>>>((const uint8_t *)W, (const uint8_t *)xq, out, M, N, nb);
    CU_LAUNCH_CHECK();
    return 0;
}

// ------------------------------------------------------------------ entry
static inline size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

// tensor-core prefill path (gemm_tc.cu)
size_t gemm_tc_workspace_bytes(int M, int K);
bool gemm_tc_supported(int wtype, int M, int N, int K);
int gemm_tc_quant_kind();
int launch_gemm_q4_0_tc(const void *W, const float *x, int ldx, const float *bias, void *y, int M, int N, int K, int ldy, int out_f16,
                        void *ws, int skip_quant, cudaStream_t st);

// k-quant weights (kquants.cu)
size_t kq_workspace_bytes(int M, int K);
int launch_linear_kq(int wtype, const void *W, const float *x, const float *bias, void *y, int M, int N, int K, int ldx, int ldy, int out_f16,
                     void *ws, int skip_quant, cudaStream_t st);

MLLMCU_API size_t mllmcu_linear_workspace_bytes(int M, int K) {
    if (M <= 0 || K <= 0) return 0;
    size_t a = align_up((size_t)M * K, 256) + align_up((size_t)M * (K / QK) * sizeof(float), 256);
    const size_t b = gemm_tc_workspace_bytes(M, K), c = kq_workspace_bytes(M, K);
    if (b > a) a = b;
    return c > a ? c : a;
}

template <int WT>
static int launch_gemv(int M, const uint8_t *W, const int8_t *xq, const float *xd, const float *bias, void *y, int N, int K,
                       int ldy, int out_f16, cudaStream_t st) {
    const dim3 grid((N + 3) / 4), block(128);
    switch (M) {
#define CASE(m) case m: k_gemv<m, WT><<<grid, block, 0, st>>>(W, xq, xd, bias, y, N, K, ldy, out_f16); break;
        CASE(1) CASE(2) CASE(3) CASE(4) CASE(5) CASE(6) CASE(7) CASE(8)
#undef CASE
    default: return -1;
    }
    CU_LAUNCH_CHECK();
    return 0;
}

int launch_gemv_mma(const mllmcu_gemv_desc &D, void *stream); // gemv_mma.cu
bool gemv_mma_supported(int wtype, int M, int K, const int *Ns, int nW, const void *const *W);
static bool linear_mma_enabled() {
    static int on = -1;
    if (on < 0) { const char *e = getenv("MLLMCU_GEMV_MMA"); on = (e && e[0] == '0') ? 0 : 1; }
    return on == 1;
}

bool gemm_mma_q4k_supported(int M, int N, int K, const void *W); // gemm_mma_q4k.cu
MLLMCU_API int mllmcu_linear_quant_kind(int wtype, int M, int N, int K, const void *W) {
    if (wtype == 12 || wtype == 14) { // 4: native block_q8_K rows at the start of ws (shared by Q4_K and Q6_K); 5: the fragment-major image of the Q4_K tensor-core kernel (kquants.cu)
        if (K % 256) return 0;
        return (wtype == 12 && gemm_mma_q4k_supported(M, N, K, W)) ? 5 : 4;
    }
    if (wtype != 2 && wtype != 8) return 0;
    if (wtype == 2 && M >= 3 && M <= 64 && linear_mma_enabled() && gemv_mma_supported(wtype, M, K, &N, 1, &W)) return 0; // own image, not in ws
    return (gemm_tc_supported(wtype, M, N, K) && !((uintptr_t)W & 15)) ? gemm_tc_quant_kind() : 3; // 3: split Q8_0 (GEMV / dp4a / scaled tcgen05)
}

MLLMCU_API int mllmcu_linear_ex(int wtype, const void *W, const float *x, const float *bias, void *y, int M, int N, int K,
                                int ldx, int ldy, int out_f16, void *ws, size_t ws_bytes, int flags, void *stream);
MLLMCU_API int mllmcu_linear(int wtype, const void *W, const float *x, const float *bias, void *y, int M, int N, int K,
                             int ldx, int ldy, int out_f16, void *ws, size_t ws_bytes, void *stream) {
    return mllmcu_linear_ex(wtype, W, x, bias, y, M, N, K, ldx, ldy, out_f16, ws, ws_bytes, 0, stream);
}

MLLMCU_API int mllmcu_linear_ex(int wtype, const void *W, const float *x, const float *bias, void *y, int M, int N, int K,
                                int ldx, int ldy, int out_f16, void *ws, size_t ws_bytes, int flags, void *stream) {
    if (M <= 0 || N <= 0 || K <= 0 || K % QK) return -1;
    if (wtype != 2 && wtype != 8 && wtype != 12 && wtype != 14) return -2;
    if (ws == nullptr || ws_bytes < mllmcu_linear_workspace_bytes(M, K)) return -3;
    cudaStream_t st = (cudaStream_t)stream;
    if (wtype == 12 || wtype == 14) return launch_linear_kq(wtype, W, x, bias, y, M, N, K, ldx, ldy, out_f16, ws, flags & 1, st);
    // 3..64 rows of Q4_0 (batched decode, short prompts): the tensor-core GEMV (gemv_mma.cu), which keeps its own activation image
    if (wtype == 2 && M >= 3 && M <= 64 && linear_mma_enabled()) {
        mllmcu_gemv_desc d = {};
        d.wtype = 2; d.nW = 1; d.W[0] = W; d.N[0] = N; d.bias[0] = bias; d.x = x; d.ldx = ldx; d.M = M; d.K = K;
        d.y = y; d.ldy = ldy; d.out_f16 = out_f16; d.x_reps = 1; d.y_reps = 1;
        const int rc = launch_gemv_mma(d, stream);
        if (rc != -2) return rc;
    }
    if (gemm_tc_supported(wtype, M, N, K) && !((uintptr_t)W & 15) && !((uintptr_t)x & 15) && !(ldx & 3)) // 16-byte loads of the activation rows
        return launch_gemm_q4_0_tc(W, x, ldx, bias, y, M, N, K, ldy, out_f16, ws, flags & 1, st);
    int8_t *xq = (int8_t *)ws;
    float *xd = (float *)((char *)ws + align_up((size_t)M * K, 256));
    if (!(flags & 1)) {
        const int rc = launch_quantize_q8_0_split(x, xq, xd, M, K, ldx, st);
        if (rc) return rc;
    }
    if (M <= GEMV_MAX_M) {
        return wtype == 2 ? launch_gemv<2>(M, (const uint8_t *)W, xq, xd, bias, y, N, K, ldy, out_f16, st)
                          : launch_gemv<8>(M, (const uint8_t *)W, xq, xd, bias, y, N, K, ldy, out_f16, st);
    }
    const dim3 grid((N + 63) / 64, (M + 63) / 64), block(256);
    if (wtype == 2) k_gemm<2><<<grid, block, 0, st>>>((const uint8_t *)W, xq, xd, bias, y, M, N, K, ldy, out_f16);
    else k_gemm<8><<<grid, block, 0, st>>>((const uint8_t *)W, xq, xd, bias, y, M, N, K, ldy, out_f16);
    CU_LAUNCH_CHECK();
    return 0;
}

// ------------------------------------------------------------------ allocation-free Linear
// The same Linear for callers that cannot allocate (CUDA graph capture, serving loops, harnesses that forbid cudaMalloc):
// the scratch is this file-scope device array, held by every GPU that loads the library until exit.  In host code the
// array's name is only a stand-in for the symbol; its device address is asked for once per device, with that device
// current, and cached.  Every call on a device shares the array (calls that overlap on the GPU race on it); nothing in it
// has to start at zero -- the quantiser writes every byte the GEMV/GEMM kernels read.
static __device__ __align__(256) unsigned char g_linear_static_ws[MLLMCU_LINEAR_STATIC_WS_BYTES];
static void *g_linear_static_ws_addr[64] = {};

MLLMCU_API size_t mllmcu_linear_static_ws_capacity(void) { return (size_t)MLLMCU_LINEAR_STATIC_WS_BYTES; }

MLLMCU_API int mllmcu_linear_static_ws(int wtype, const void *W, const float *x, const float *bias, void *y, int M, int N, int K,
                                       int ldx, int ldy, int out_f16, void *stream) {
    if (M <= 0 || N <= 0 || K <= 0 || K % QK) return -1;
    if (mllmcu_linear_workspace_bytes(M, K) > (size_t)MLLMCU_LINEAR_STATIC_WS_BYTES) {
        fprintf(stderr, "[mllmcu] mllmcu_linear_static_ws: M=%d K=%d needs %zu bytes of scratch, the static array holds %zu\n", M, K,
                mllmcu_linear_workspace_bytes(M, K), (size_t)MLLMCU_LINEAR_STATIC_WS_BYTES);
        return -3;
    }
    int dev = -1;
    CU_TRY(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64) return -1;
    void *ws = __atomic_load_n(&g_linear_static_ws_addr[dev], __ATOMIC_ACQUIRE);
    if (ws == nullptr) { // once per device; racing first calls resolve the same address
        CU_TRY(cudaGetSymbolAddress(&ws, g_linear_static_ws));
        if (ws == nullptr) return -1;
        __atomic_store_n(&g_linear_static_ws_addr[dev], ws, __ATOMIC_RELEASE);
    }
    return mllmcu_linear_ex(wtype, W, x, bias, y, M, N, K, ldx, ldy, out_f16, ws, (size_t)MLLMCU_LINEAR_STATIC_WS_BYTES, 0, stream);
}
