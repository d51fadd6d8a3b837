// kquants.cu -- k-quant Linear: Q4_K / Q6_K weights x Q8_K activations.
// Reference: mat_mul generic path (compute/Matmul.cpp:104-115 row quantisation to vec_dot_type Q8_K, :223-276 vec_dot
// loop) with quantize_row_q8_K (ggml/QuantizeQ8.cpp:216-266), vec_dot_q4_K_q8_K (ggml/VecDotQ4.cpp:32-345, AVX2 body
// :220-283) and vec_dot_q6_K_q8_K (ggml/VecDotQ6.cpp:32-262).
//
// Weights stay in the .mllm file's native super-blocks (block_q4_K 144 B, block_q6_K 210 B per 256 weights,
// mllm/DataType.hpp:93-98, :123-128).  Activations are quantised once per call into native block_q8_K (292 B: float d,
// 256 x int8, 16 x int16 group sums), bytes identical to the reference's.  Per super-block the kernels form the same
// two exact integers the reference forms -- sum_sb scale_sb * (q . q8) and sum_sb min_sb * bsum_sb -- and apply
// d = d_x * fp16(d_w), dmin = -d_x * fp16(dmin_w) with one fp32 FMA each; only the order of the fp32 accumulation over
// super-blocks differs from the 8 AVX lanes (tolerance in the tests).
#include "common.cuh"
#include <algorithm>
#include <stdlib.h>
#include "kq_common.cuh"

#define QK_K 256
#define Q4K_BYTES 144
#define Q6K_BYTES 210
#define Q8K_BYTES 292
#define Q8K_PITCH 304 // the decode chain's re-pitched block: d at 0, qs at 16 (16-byte aligned), the 16 group sums at 272

// nearest_int (ggml/Quantize.hpp:174-180)
static __device__ __forceinline__ int nearest_int_dev(float f) {
    const float v = __fadd_rn(f, 12582912.f);
    return (int)(__float_as_uint(v) & 0x007fffffu) - 0x00400000;
}

// ------------------------------------------------------------ quantize_row_q8_K
// one warp per super-block, lane l owns elements 8l..8l+7 (v) and writes its share of the 292-byte block at `out`
// the warp's super-block quantised: w0 / w1 = this lane's eight int8 (elements 8 lane .. 8 lane + 7), s2 = the sum of the 16-element group the
// lane pair (lane, lane ^ 1) covers, d = the block scale
static __device__ __forceinline__ void q8K_quant_warp(const float (&v)[8], int lane, uint32_t &w0, uint32_t &w1, int &s2, float &d) {
    // first element with the largest |x| decides the sign of the scale (strict > in the reference's scan)
    float amax = 0.f, mx = 0.f;
    int idx = 0x7fffffff;
#pragma unroll
    for (int j = 0; j < 8; j++) {
        const float ax = fabsf(v[j]);
        if (ax > amax) { amax = ax; mx = v[j]; idx = lane * 8 + j; }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const float oa = __shfl_xor_sync(0xffffffffu, amax, o);
        const float om = __shfl_xor_sync(0xffffffffu, mx, o);
        const int oi = __shfl_xor_sync(0xffffffffu, idx, o);
        if (oa > amax || (oa == amax && oi < idx)) { amax = oa; mx = om; idx = oi; }
    }
    int q[8];
    d = 0.f;
    if (amax == 0.f) {
#pragma unroll
        for (int j = 0; j < 8; j++) q[j] = 0;
    } else {
        const float iscale = __fdiv_rn(-128.f, mx);
#pragma unroll
        for (int j = 0; j < 8; j++) q[j] = min(127, nearest_int_dev(__fmul_rn(iscale, v[j])));
        d = __fdiv_rn(1.f, iscale);
    }
    int s = 0;
#pragma unroll
    for (int j = 0; j < 8; j++) s += q[j];
    s2 = s + __shfl_xor_sync(0xffffffffu, s, 1);
    w0 = 0; w1 = 0;
#pragma unroll
    for (int j = 0; j < 4; j++) { w0 |= (uint32_t)(q[j] & 0xFF) << (8 * j); w1 |= (uint32_t)(q[4 + j] & 0xFF) << (8 * j); }
}

// one warp per super-block, lane l owns elements 8l..8l+7 (v) and writes its share of the 292-byte block at `out`
static __device__ __forceinline__ void q8K_block_from_warp(const float (&v)[8], int lane, uint8_t *out, int qs_off = 4, int bs_off = 260) {
    uint32_t w0, w1;
    int s2;
    float d;
    q8K_quant_warp(v, lane, w0, w1, s2, d);
    *(uint32_t *)(out + qs_off + lane * 8) = w0; // blocks are 292 B apart: only 4-byte alignment is guaranteed
    *(uint32_t *)(out + qs_off + 4 + lane * 8) = w1;
    if ((lane & 1) == 0) *(int16_t *)(out + bs_off + (lane >> 1) * 2) = (int16_t)s2; // all-zero block: 0 (the reference leaves it unwritten)
    if (lane == 0) *(float *)out = d;
}

// The activation image of the Q4_K tensor-core Linear (gemm_mma_q4k.cu), "fragment-major": the same block_q8_K values, placed so that
// every operand load of a warp is contiguous.  Rows in groups of 8 (G groups, the last one padded), nb super-blocks:
//   QS  [G][nb][8 sub-blocks][32 lanes = row-in-group * 4 + t][8 int8: elements 8t..8t+7 of the sub-block]   2048 B per (group, super-block)
//   DX  [G][nb][8 rows] float d
//   BS  [G][nb][8 rows][8] int16: bsums[2j] + bsums[2j+1], the sum of sub-block j
// A B fragment of mma.m16n8k32 (8 activation rows x 32 elements, lane (g, t) = row g, elements 8t..8t+7 with the relabelled contraction
// index) is then 256 contiguous bytes instead of 8 rows in 8 cache lines.
__global__ void __launch_bounds__(128) k_quantize_q8_K_frag(const float *__restrict__ x, uint8_t *__restrict__ y, long long nblk, int nb, int ldx, int G) {
    const int lane = threadIdx.x & 31;
    const long long blk = (long long)blockIdx.x * 4 + (threadIdx.x >> 5);
    if (blk >= nblk) return;
    const int row = (int)(blk / nb), b = (int)(blk % nb);
    const float *xp = x + (size_t)row * ldx + (size_t)b * QK_K + lane * 8;
    const float4 v0 = *(const float4 *)xp, v1 = *(const float4 *)(xp + 4);
    const float v[8] = {v0.x, v0.y, v0.z, v0.w, v1.x, v1.y, v1.z, v1.w};
    uint32_t w0, w1;
    int s2;
    float d;
    q8K_quant_warp(v, lane, w0, w1, s2, d);
    const int grp = row >> 3, g = row & 7, j = lane >> 2, t = lane & 3;
    const size_t cell = (size_t)grp * nb + b;
    *(uint2 *)(y + cell * 2048 + j * 256 + (g * 4 + t) * 8) = make_uint2(w0, w1);
    const int s4 = s2 + __shfl_xor_sync(0xffffffffu, s2, 2); // lanes 4j..4j+3: the 32 elements of sub-block j
    uint8_t *dx = y + (size_t)G * nb * 2048, *bs = dx + (size_t)G * nb * 32;
    if (t == 0) *(int16_t *)(bs + cell * 128 + g * 16 + j * 2) = (int16_t)s4;
    if (lane == 0) *(float *)(dx + cell * 32 + g * 4) = d;
}

size_t kq_frag_image_bytes(int M, int K) { return (size_t)((M + 7) / 8) * (K / QK_K) * (2048 + 32 + 128) + 256; }
int launch_quantize_q8_K_frag(const float *x, void *y, int M, int K, int ldx, cudaStream_t st) {
    if (M <= 0 || K <= 0 || K % QK_K || ldx % 4 || ((uintptr_t)x & 15) || ((uintptr_t)y & 15)) return -1;
    const long long nblk = (long long)M * (K / QK_K);
    k_quantize_q8_K_frag<<<(unsigned)((nblk + 3) / 4), 128, 0, st>>>(x, (uint8_t *)y, nblk, K / QK_K, ldx, (M + 7) / 8);
    CU_LAUNCH_CHECK();
    return 0;
}

__global__ void __launch_bounds__(128) k_quantize_q8_K(const float *__restrict__ x, uint8_t *__restrict__ y, long long nblk, int nb, int ldx) {
    const int lane = threadIdx.x & 31;
    const long long blk = (long long)blockIdx.x * 4 + (threadIdx.x >> 5);
    if (blk >= nblk) return;
    const long long row = blk / nb;
    const int b = (int)(blk % nb);
    const float *xp = x + (size_t)row * ldx + (size_t)b * QK_K + lane * 8;
    const float4 v0 = *(const float4 *)xp, v1 = *(const float4 *)(xp + 4);
    const float v[8] = {v0.x, v0.y, v0.z, v0.w, v1.x, v1.y, v1.z, v1.w};
    q8K_block_from_warp(v, lane, y + (size_t)blk * Q8K_BYTES);
}

MLLMCU_API int mllmcu_quantize_q8_K(const float *x, void *y_blocks, int rows, int k, int ldx, void *stream) {
    if (rows <= 0 || k <= 0 || k % QK_K || ldx % 4 || ((uintptr_t)x & 15) || ((uintptr_t)y_blocks & 3)) return -1;
    const long long nblk = (long long)rows * (k / QK_K);
    k_quantize_q8_K<<<(unsigned)((nblk + 3) / 4), 128, 0, (cudaStream_t)stream>>>(x, (uint8_t *)y_blocks, nblk, k / QK_K, ldx);
    CU_LAUNCH_CHECK();
    return 0;
}

// ------------------------------------------------------------ weight decode
// get_scale_min_k4 (ggml/QuantizeQ4.cpp:177-184)
static __device__ __forceinline__ void q4k_scale_min(const uint8_t *q, int j, int &sc, int &mn) {
    if (j < 4) { sc = q[j] & 63; mn = q[j + 4] & 63; }
    else { sc = (q[j + 4] & 0xF) | ((q[j - 4] >> 6) << 4); mn = (q[j + 4] >> 4) | ((q[j] >> 6) << 4); }
}
// the 32 weights of chunk c (elements 32c..32c+31) of a super-block as 8 words of int8
//   Q4_K: values 0..15 (the min term carries the offset); Q6_K: values -32..31
template <int WT>
static __device__ __forceinline__ void kq_chunk_words(const uint8_t *blk, int c, int *w) {
    if (WT == 12) {
        const uint4 *p = (const uint4 *)(blk + 16 + (c >> 1) * 32);
        const uint4 a = p[0], b = p[1];
        const uint32_t r[8] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w};
#pragma unroll
        for (int j = 0; j < 8; j++) w[j] = (int)((c & 1) ? ((r[j] >> 4) & 0x0F0F0F0Fu) : (r[j] & 0x0F0F0F0Fu));
    } else {
        const int n = c >> 2, g = c & 3;
        const uint16_t *ql = (const uint16_t *)(blk + n * 64 + ((g & 1) ? 32 : 0));
        const uint16_t *qh = (const uint16_t *)(blk + 128 + n * 32);
#pragma unroll
        for (int j = 0; j < 8; j++) {
            const uint32_t l = (uint32_t)ql[2 * j] | ((uint32_t)ql[2 * j + 1] << 16);
            const uint32_t h = (uint32_t)qh[2 * j] | ((uint32_t)qh[2 * j + 1] << 16);
            const uint32_t nib = (g >= 2) ? ((l >> 4) & 0x0F0F0F0Fu) : (l & 0x0F0F0F0Fu);
            const uint32_t hi = ((h >> (2 * g)) & 0x03030303u) << 4;
            w[j] = (int)__vsub4(nib | hi, 0x20202020u);
        }
    }
}

// ------------------------------------------------------------------- GEMV
// M <= 8 rows (decode).  One warp per output row; lane l takes chunk l%8 of super-block l/8 + 4*it, the 8 lanes of a
// super-block add their integers (exactly what the reference sums) before the fp32 FMA.
template <int WT, int M>
__global__ void __launch_bounds__(128) k_gemv_kq(const uint8_t *__restrict__ W, const uint8_t *__restrict__ xq, const float *__restrict__ bias,
                                                 void *y, int N, int nb, int ldy, int out_f16) {
    constexpr int BS = (WT == 12) ? Q4K_BYTES : Q6K_BYTES;
    const int lane = threadIdx.x & 31;
    const int n = blockIdx.x * 4 + (threadIdx.x >> 5);
    if (n >= N) return;
    const uint8_t *row = W + (size_t)n * nb * BS;
    const int c = lane & 7;
    float acc[M], accm[M];
#pragma unroll
    for (int m = 0; m < M; m++) { acc[m] = 0.f; accm[m] = 0.f; }
    for (int i0 = 0; i0 < nb; i0 += 4) {
        const int i = i0 + (lane >> 3);
        const bool on = i < nb;
        const uint8_t *blk = row + (size_t)(on ? i : 0) * BS;
        int w[8];
        kq_chunk_words<WT>(blk, c, w);
        int sc0, sc1 = 0, mn = 0;
        float dw, dminw = 0.f;
        if (WT == 12) {
            q4k_scale_min(blk + 4, c, sc0, mn);
            dw = __half2float(*(const __half *)blk);
            dminw = __half2float(*(const __half *)(blk + 2));
        } else {
            sc0 = (int)(int8_t)blk[192 + 2 * c];
            sc1 = (int)(int8_t)blk[192 + 2 * c + 1];
            dw = __half2float(*(const __half *)(blk + 208));
        }
#pragma unroll
        for (int m = 0; m < M; m++) {
            const uint8_t *xb = xq + ((size_t)m * nb + (on ? i : 0)) * Q8K_BYTES;
            const int *a = (const int *)(xb + 4 + c * 32);
            int vi, vm = 0;
            if (WT == 12) {
                int is = 0;
#pragma unroll
                for (int j = 0; j < 8; j++) is = __dp4a(w[j], a[j], is);
                vi = sc0 * is;
                const int16_t *bs = (const int16_t *)(xb + 260);
                vm = mn * ((int)bs[2 * c] + (int)bs[2 * c + 1]);
            } else {
                int i0s = 0, i1s = 0;
#pragma unroll
                for (int j = 0; j < 4; j++) { i0s = __dp4a(w[j], a[j], i0s); i1s = __dp4a(w[4 + j], a[4 + j], i1s); }
                vi = sc0 * i0s + sc1 * i1s;
            }
            if (!on) { vi = 0; vm = 0; }
#pragma unroll
            for (int o = 1; o < 8; o <<= 1) {
                vi += __shfl_xor_sync(0xffffffffu, vi, o);
                if (WT == 12) vm += __shfl_xor_sync(0xffffffffu, vm, o);
            }
            if (c == 0 && on) {
                const float dx = *(const float *)xb;
                acc[m] = fmaf(__fmul_rn(dx, dw), (float)vi, acc[m]);
                if (WT == 12) accm[m] = fmaf(__fmul_rn(-dx, dminw), (float)vm, accm[m]);
            }
        }
    }
#pragma unroll
    for (int m = 0; m < M; m++) {
        float v = warp_sum(acc[m]);
        if (WT == 12) v = __fadd_rn(v, warp_sum(accm[m]));
        if (lane == 0) {
            if (bias) v = __fadd_rn(v, bias[n]);
            if (out_f16) ((__half *)y)[(size_t)m * ldy + n] = __float2half_rn(v);
            else ((float *)y)[(size_t)m * ldy + n] = v;
        }
    }
}

// ------------------------------------------------------------------- GEMM
// 64x64 output tile per CTA of 256 threads (4x4 per thread); one 32-element chunk per step, integer accumulators carried
// across the 8 chunks of a super-block and folded into fp32 once per super-block.
template <int WT>
__global__ void __launch_bounds__(256) k_gemm_kq(const uint8_t *__restrict__ W, const uint8_t *__restrict__ xq, const float *__restrict__ bias,
                                                 void *y, int M, int N, int nb, int ldy, int out_f16) {
    constexpr int BM = 64, BN = 64;
    constexpr int BS = (WT == 12) ? Q4K_BYTES : Q6K_BYTES;
    __shared__ int sA[BM][9];
    __shared__ int sB[BN][9];
    __shared__ int sAbs[BM];           // bsums pair of the chunk
    __shared__ int sBsc[BN][2], sBmn[BN];
    __shared__ float sAd[BM], sBd[BN], sBdm[BN];
    const int m0 = blockIdx.y * BM, n0 = blockIdx.x * BN;
    const int t = threadIdx.x, tx = t & 15, ty = t >> 4;
    float acc[4][4], accm[4][4];
    int iacc[4][4], macc[4][4];
#pragma unroll
    for (int i = 0; i < 4; i++)
#pragma unroll
        for (int j = 0; j < 4; j++) { acc[i][j] = 0.f; accm[i][j] = 0.f; iacc[i][j] = 0; macc[i][j] = 0; }
    for (int sbk = 0; sbk < nb; sbk++) {
        for (int c = 0; c < 8; c++) {
            if (t < 128) { // activations: 64 rows x 2 halves of 4 words
                const int r = t >> 1, h = t & 1;
                int4 v = make_int4(0, 0, 0, 0);
                if (m0 + r < M) {
                    const uint8_t *xb = xq + ((size_t)(m0 + r) * nb + sbk) * Q8K_BYTES;
                    const int *a = (const int *)(xb + 4 + c * 32 + h * 16);
                    v = make_int4(a[0], a[1], a[2], a[3]);
                    if (h == 0) {
                        const int16_t *bs = (const int16_t *)(xb + 260);
                        sAbs[r] = (int)bs[2 * c] + (int)bs[2 * c + 1];
                        if (c == 0) sAd[r] = *(const float *)xb;
                    }
                } else if (h == 0) { sAbs[r] = 0; if (c == 0) sAd[r] = 0.f; }
                sA[r][h * 4 + 0] = v.x; sA[r][h * 4 + 1] = v.y; sA[r][h * 4 + 2] = v.z; sA[r][h * 4 + 3] = v.w;
            } else if (t < 192) { // weights: one row's chunk per thread
                const int r = t - 128;
                int w[8] = {0, 0, 0, 0, 0, 0, 0, 0};
                int sc0 = 0, sc1 = 0, mn = 0;
                if (n0 + r < N) {
                    const uint8_t *blk = W + ((size_t)(n0 + r) * nb + sbk) * BS;
                    kq_chunk_words<WT>(blk, c, w);
                    if (WT == 12) {
                        q4k_scale_min(blk + 4, c, sc0, mn);
                        sc1 = sc0;
                        if (c == 0) { sBd[r] = __half2float(*(const __half *)blk); sBdm[r] = __half2float(*(const __half *)(blk + 2)); }
                    } else {
                        sc0 = (int)(int8_t)blk[192 + 2 * c];
                        sc1 = (int)(int8_t)blk[192 + 2 * c + 1];
                        if (c == 0) { sBd[r] = __half2float(*(const __half *)(blk + 208)); sBdm[r] = 0.f; }
                    }
                } else if (c == 0) { sBd[r] = 0.f; sBdm[r] = 0.f; }
#pragma unroll
                for (int j = 0; j < 8; j++) sB[r][j] = w[j];
                sBsc[r][0] = sc0; sBsc[r][1] = sc1; sBmn[r] = mn;
            }
            __syncthreads();
            int lo[4][4], hi[4][4];
#pragma unroll
            for (int i = 0; i < 4; i++)
#pragma unroll
                for (int j = 0; j < 4; j++) { lo[i][j] = 0; hi[i][j] = 0; }
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
                    for (int j = 0; j < 4; j++) {
                        if (w < 4) lo[i][j] = __dp4a(a[i], bb[j], lo[i][j]);
                        else hi[i][j] = __dp4a(a[i], bb[j], hi[i][j]);
                    }
            }
#pragma unroll
            for (int j = 0; j < 4; j++) {
                const int s0 = sBsc[tx * 4 + j][0], s1 = sBsc[tx * 4 + j][1], mn = sBmn[tx * 4 + j];
#pragma unroll
                for (int i = 0; i < 4; i++) {
                    iacc[i][j] += s0 * lo[i][j] + s1 * hi[i][j];
                    if (WT == 12) macc[i][j] += mn * sAbs[ty * 4 + i];
                }
            }
            if (c == 7) {
#pragma unroll
                for (int i = 0; i < 4; i++)
#pragma unroll
                    for (int j = 0; j < 4; j++) {
                        const float dx = sAd[ty * 4 + i];
                        acc[i][j] = fmaf(__fmul_rn(dx, sBd[tx * 4 + j]), (float)iacc[i][j], acc[i][j]);
                        if (WT == 12) accm[i][j] = fmaf(__fmul_rn(-dx, sBdm[tx * 4 + j]), (float)macc[i][j], accm[i][j]);
                        iacc[i][j] = 0; macc[i][j] = 0;
                    }
            }
            __syncthreads();
        }
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
            if (WT == 12) v = __fadd_rn(v, accm[i][j]);
            if (bias) v = __fadd_rn(v, bias[n]);
            if (out_f16) ((__half *)y)[(size_t)m * ldy + n] = __float2half_rn(v);
            else ((float *)y)[(size_t)m * ldy + n] = v;
        }
    }
}

size_t kq_workspace_bytes(int M, int K) { return (K % QK_K) ? 0 : std::max((size_t)M * (K / QK_K) * Q8K_PITCH + 256, kq_frag_image_bytes(M, K)); } // room for every image form

template <int WT>
static int launch_kq(const uint8_t *W, const uint8_t *xq, const float *bias, void *y, int M, int N, int nb, int ldy, int out_f16, cudaStream_t st) {
    if (M <= 8) {
        const dim3 grid((N + 3) / 4), block(128);
        switch (M) {
#define CASE(m) case m: k_gemv_kq<WT, m><<<grid, block, 0, st>>>(W, xq, bias, y, N, nb, ldy, out_f16); break;
            CASE(1) CASE(2) CASE(3) CASE(4) CASE(5) CASE(6) CASE(7) CASE(8)
#undef CASE
        }
    } else {
        const dim3 grid((N + 63) / 64, (M + 63) / 64);
        k_gemm_kq<WT><<<grid, 256, 0, st>>>(W, xq, bias, y, M, N, nb, ldy, out_f16);
    }
    CU_LAUNCH_CHECK();
    return 0;
}

int launch_gemm_mma_q4k(const void *W, const void *xq, const float *bias, void *y, int M, int N, int K, int ldy, int out_f16, cudaStream_t st); // gemm_mma_q4k.cu
bool gemm_mma_q4k_supported(int M, int N, int K, const void *W);
MLLMCU_API int mllmcu_prep_q8_K(int mode, const float *x, const float *aux, float eps, int add_unit_offset, void *y_blocks, int rows, int k, int ldx,
                                int flags, void *stream); // below
// called by mllmcu_linear for wtype 12 (Q4_K) / 14 (Q6_K)
int launch_linear_kq(int wtype, const void *W, const float *x, const float *bias, void *y, int M, int N, int K, int ldx, int ldy, int out_f16,
                     void *ws, int skip_quant, cudaStream_t st) {
    if (K % QK_K || ldx % 4 || ((uintptr_t)x & 15)) return -1;
    if (wtype == 12 && ((uintptr_t)W & 15)) return -1;
    if (wtype == 14 && ((uintptr_t)W & 1)) return -1;
    if (wtype == 12 && gemm_mma_q4k_supported(M, N, K, W)) { // 9+ rows: the tensor-core kernel (gemm_mma_q4k.cu) on the fragment-major image
        if (!skip_quant) { // skip: ws already holds this image of these activations (q / k / v, gate / up share one input)
            const int rc = launch_quantize_q8_K_frag(x, ws, M, K, ldx, st);
            if (rc) return rc;
        }
        const int rm = launch_gemm_mma_q4k(W, ws, bias, y, M, N, K, ldy, out_f16, st);
        if (rm != -2) return rm;
        skip_quant = 0; // a shape the kernel turned down after all: the kernels below read native blocks
    }
    if (!skip_quant) {
        const int rc = mllmcu_quantize_q8_K(x, ws, M, K, ldx, st);
        if (rc) return rc;
    }
    return wtype == 12 ? launch_kq<12>((const uint8_t *)W, (const uint8_t *)ws, bias, y, M, N, K / QK_K, ldy, out_f16, st)
                       : launch_kq<14>((const uint8_t *)W, (const uint8_t *)ws, bias, y, M, N, K / QK_K, ldy, out_f16, st);
}

// ------------------------------------------------------------------- decode-step chain for Q4_K bodies
// The single-token step of a Q4_K model (LLaVA-1.5-7B, BASELINE configs[2]) as nine launches per decoder block instead of eighteen:
//   prep(RMSNorm) -> GEMV[q|k|v] -> attention -> prep -> GEMV[o] + residual -> prep(RMSNorm) -> GEMV[gate|up] -> prep(SiLU * up) -> GEMV[down] + residual
// "prep" writes the block_q8_K image of the NEXT Linear's input, folding the op in front of it (CPURMSNorm.cpp:31-95, CPUSiLU / CPUMul)
// into the quantiser with the arithmetic of k_rmsnorm / k_ew, so every value is the one the op-by-op sequence produces.  The GEMV reads
// up to three matrices that share the activation rows and can add the residual stream in its epilogue.  Every kernel is a programmatic
// dependent of the one before it: it fetches what does not depend on its predecessor (weights) before griddepcontrol.wait.
static __device__ __forceinline__ void kq_pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
static __device__ __forceinline__ void kq_pdl_launch() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }

// grid (ceil(nb / 8), rows), 8 warps: warp w of CTA c owns super-block 8c + w of its row.  MODE 0: x; 1: RMSNorm(x; aux = weight);
// 2: SiLU(x) * aux (aux has x's layout).  With MODE 1 every CTA of a row repeats k_rmsnorm's sum (same order, same rounding).
template <int MODE>
__global__ void __launch_bounds__(256) k_prep_q8_K(const float *__restrict__ x, const float *__restrict__ aux, uint8_t *__restrict__ y, int nb,
                                                   int ldx, float eps, int add_unit_offset, int aligned) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int row = blockIdx.y, b = blockIdx.x * 8 + warp;
    const float *xr = x + (size_t)row * ldx;
    kq_pdl_launch();
    kq_pdl_wait();
    float rms = 1.f;
    if (MODE == 1) {
        const int dim = nb * QK_K;
        const float4 *x4 = (const float4 *)xr;
        double s0 = 0.0, s1 = 0.0;
        for (int d = threadIdx.x; d < (dim >> 2); d += 256) {
            const float4 v = x4[d];
            s0 = fma((double)v.x, (double)v.x, s0); s1 = fma((double)v.y, (double)v.y, s1);
            s0 = fma((double)v.z, (double)v.z, s0); s1 = fma((double)v.w, (double)v.w, s1);
        }
        __shared__ double sred[8];
        __shared__ float s_rms;
        const double ss = warp_sum_d(s0 + s1);
        if (lane == 0) sred[warp] = ss;
        __syncthreads();
        if (threadIdx.x == 0) {
            double t = 0.0;
            for (int i = 0; i < 8; i++) t += sred[i];
            const float mean = (float)(t / (double)dim);
            s_rms = __fdiv_rn(1.0f, __fsqrt_rn(__fadd_rn(mean, eps)));
        }
        __syncthreads();
        rms = s_rms;
    }
    if (b >= nb) return;
    const float *xp = xr + (size_t)b * QK_K + lane * 8;
    const float4 v0 = *(const float4 *)xp, v1 = *(const float4 *)(xp + 4);
    float v[8] = {v0.x, v0.y, v0.z, v0.w, v1.x, v1.y, v1.z, v1.w};
    if (MODE == 1) {
        const float4 w0 = *(const float4 *)(aux + (size_t)b * QK_K + lane * 8), w1 = *(const float4 *)(aux + (size_t)b * QK_K + lane * 8 + 4);
        float w[8] = {w0.x, w0.y, w0.z, w0.w, w1.x, w1.y, w1.z, w1.w};
#pragma unroll
        for (int j = 0; j < 8; j++) {
            if (add_unit_offset) w[j] = __fadd_rn(1.0f, w[j]);
            v[j] = __fmul_rn(__fmul_rn(v[j], rms), w[j]);
        }
    } else if (MODE == 2) {
        const float *up = aux + (size_t)row * ldx + (size_t)b * QK_K + lane * 8;
        const float4 u0 = *(const float4 *)up, u1 = *(const float4 *)(up + 4);
        const float u[8] = {u0.x, u0.y, u0.z, u0.w, u1.x, u1.y, u1.z, u1.w};
#pragma unroll
        for (int j = 0; j < 8; j++) v[j] = __fmul_rn(__fdiv_rn(v[j], __fadd_rn(1.0f, v_expf(__fsub_rn(0.0f, v[j])))), u[j]); // silu_f (elementwise.cu)
    }
    if (aligned) q8K_block_from_warp(v, lane, y + ((size_t)row * nb + b) * Q8K_PITCH, 16, 272);
    else q8K_block_from_warp(v, lane, y + ((size_t)row * nb + b) * Q8K_BYTES);
}

static cudaError_t kq_launch_cfg(cudaLaunchConfig_t &cfg, cudaLaunchAttribute *attr, dim3 grid, dim3 block, int pdl, cudaStream_t st) {
    cfg = cudaLaunchConfig_t{};
    cfg.gridDim = grid; cfg.blockDim = block; cfg.stream = st;
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = pdl ? 1 : 0;
    return cudaSuccess;
}

MLLMCU_API int mllmcu_prep_q8_K(int mode, const float *x, const float *aux, float eps, int add_unit_offset, void *y_blocks, int rows, int k, int ldx,
                                int flags, void *stream) {
    const int pdl = flags & 1, aligned = (flags & 2) ? 1 : 0;
    if (aligned && ((uintptr_t)y_blocks & 15)) return -1;
    if (mode < 0 || mode > 2 || rows <= 0 || k <= 0 || k % QK_K || ldx % 4 || ((uintptr_t)x & 15) || ((uintptr_t)y_blocks & 3)) return -1;
    if (mode != 0 && (aux == nullptr || ((uintptr_t)aux & 15))) return -1;
    const int nb = k / QK_K;
    cudaLaunchConfig_t cfg;
    cudaLaunchAttribute attr[1];
    kq_launch_cfg(cfg, attr, dim3((unsigned)((nb + 7) / 8), (unsigned)rows), dim3(256), pdl, (cudaStream_t)stream);
    uint8_t *y = (uint8_t *)y_blocks;
    if (mode == 0) CU_TRY(cudaLaunchKernelEx(&cfg, k_prep_q8_K<0>, x, aux, y, nb, ldx, eps, add_unit_offset, aligned));
    else if (mode == 1) CU_TRY(cudaLaunchKernelEx(&cfg, k_prep_q8_K<1>, x, aux, y, nb, ldx, eps, add_unit_offset, aligned));
    else CU_TRY(cudaLaunchKernelEx(&cfg, k_prep_q8_K<2>, x, aux, y, nb, ldx, eps, add_unit_offset, aligned));
    return 0;
}

struct KqGemvParams {
    const uint8_t *W[3];
    const float *bias[3];
    float *y[3];
    int N[3];
    int nW, Ntot, nb, ldy, ldr;
    const uint8_t *xq;
    const float *res;
};

// One warp per output row, M <= 8 activation rows.  Lane (g, t) = (lane / 8, lane % 8) takes 16 weight bytes of super-block g + 4 it:
// bytes [16 (t & 1), +16) of nibble pair t / 2, i.e. half of sub-block 2 (t / 2) in the low nibbles and half of sub-block 2 (t / 2) + 1 in
// the high ones -- every weight byte is fetched by exactly one lane, 16 bytes at a time, and four super-block groups (up to 16 super-
// blocks, 2.3 KB per warp) are in flight before the first is consumed.  Each lane keeps its own exact integers (scale * dot and
// min * group sum over its 32 weights) and folds them into fp32 accumulators of its own with the super-block's two factors; the lanes meet
// once per row.  Against the reference (integers summed over the whole super-block before the FMA pair) only the fp32 summation order
// differs, as it already does between its AVX lanes and any other order (tolerance in the tests).  Measured (profiles/r02_q4k_gemv_micro_*):
// the kernel is issue-bound, not byte-bound, which is why the per-super-block shuffles and the unaligned activation loads were removed.
// ALIGNED: the activation blocks are the 304-byte re-pitched form (qs 16-byte aligned), else native block_q8_K (4-byte aligned).
template <int M, bool ALIGNED>
__global__ void __launch_bounds__(128) k_gemv_q4k(const KqGemvParams p) {
    constexpr int XB = ALIGNED ? Q8K_PITCH : Q8K_BYTES, XQS = ALIGNED ? 16 : 4, XBS = ALIGNED ? 272 : 260;
    const int lane = threadIdx.x & 31;
    int n = blockIdx.x * 4 + (threadIdx.x >> 5);
    if (n >= p.Ntot) return;
    int s = 0;
    if (n >= p.N[0]) { n -= p.N[0]; s = 1; if (n >= p.N[1]) { n -= p.N[1]; s = 2; } }
    const int nb = p.nb, g = lane >> 3, t = lane & 7, j = t >> 1, h = t & 1;
    // selects, not p.W[s]: a run-time index would copy the parameter arrays to local memory
    const uint8_t *row = (s == 0 ? p.W[0] : s == 1 ? p.W[1] : p.W[2]) + (size_t)n * nb * Q4K_BYTES;
    float acc[M], accm[M];
#pragma unroll
    for (int m = 0; m < M; m++) { acc[m] = 0.f; accm[m] = 0.f; }
    for (int i0 = 0; i0 < nb; i0 += 16) {
        uint4 hq[4], wq[4];
#pragma unroll
        for (int u = 0; u < 4; u++) {
            const int i = i0 + 4 * u + g;
            const uint8_t *blk = row + (size_t)(i < nb ? i : 0) * Q4K_BYTES;
            hq[u] = ld_stream_u4(blk);                       // d, dmin, the 12 scale bytes
            wq[u] = ld_stream_u4(blk + 16 + t * 16);
        }
        if (i0 == 0) { kq_pdl_launch(); kq_pdl_wait(); }     // the activations are the predecessor's output
#pragma unroll
        for (int u = 0; u < 4; u++) {
            const int i = i0 + 4 * u + g;
            if (i >= nb) continue;
            uint32_t sc[2], mn[2];
            q4k_scales8(hq[u].y, hq[u].z, hq[u].w, sc, mn);
            const int sc_lo = byte_of(sc, 2 * j), sc_hi = byte_of(sc, 2 * j + 1), mn_lo = byte_of(mn, 2 * j), mn_hi = byte_of(mn, 2 * j + 1);
            const float dw = __half2float(__ushort_as_half((unsigned short)(hq[u].x & 0xFFFFu)));
            const float dminw = __half2float(__ushort_as_half((unsigned short)(hq[u].x >> 16)));
            const uint32_t wl[4] = {wq[u].x & 0x0F0F0F0Fu, wq[u].y & 0x0F0F0F0Fu, wq[u].z & 0x0F0F0F0Fu, wq[u].w & 0x0F0F0F0Fu};
            const uint32_t wh[4] = {(wq[u].x >> 4) & 0x0F0F0F0Fu, (wq[u].y >> 4) & 0x0F0F0F0Fu, (wq[u].z >> 4) & 0x0F0F0F0Fu, (wq[u].w >> 4) & 0x0F0F0F0Fu};
#pragma unroll
            for (int m = 0; m < M; m++) {
                const uint8_t *xb = p.xq + ((size_t)m * nb + i) * XB;
                int al[4], ah[4];
                if (ALIGNED) {
                    const int4 a0 = *(const int4 *)(xb + XQS + (2 * j) * 32 + h * 16), a1 = *(const int4 *)(xb + XQS + (2 * j + 1) * 32 + h * 16);
                    al[0] = a0.x; al[1] = a0.y; al[2] = a0.z; al[3] = a0.w; ah[0] = a1.x; ah[1] = a1.y; ah[2] = a1.z; ah[3] = a1.w;
                } else {
                    const int *pl = (const int *)(xb + XQS + (2 * j) * 32 + h * 16), *ph = (const int *)(xb + XQS + (2 * j + 1) * 32 + h * 16);
#pragma unroll
                    for (int q = 0; q < 4; q++) { al[q] = pl[q]; ah[q] = ph[q]; }
                }
                int is_lo = 0, is_hi = 0;
#pragma unroll
                for (int q = 0; q < 4; q++) {
                    is_lo = __dp4a((int)wl[q], al[q], is_lo);
                    is_hi = __dp4a((int)wh[q], ah[q], is_hi);
                }
                const int16_t *bs = (const int16_t *)(xb + XBS);
                const int vi = sc_lo * is_lo + sc_hi * is_hi;                                   // |vi| < 2^23: exact in fp32
                const int vm = mn_lo * (int)bs[4 * j + h] + mn_hi * (int)bs[4 * j + 2 + h];
                const float dx = *(const float *)xb;
                acc[m] = fmaf(__fmul_rn(dx, dw), (float)vi, acc[m]);
                accm[m] = fmaf(__fmul_rn(-dx, dminw), (float)vm, accm[m]);
            }
        }
    }
    const float *bias = s == 0 ? p.bias[0] : s == 1 ? p.bias[1] : p.bias[2];
    float *y = s == 0 ? p.y[0] : s == 1 ? p.y[1] : p.y[2];
#pragma unroll
    for (int m = 0; m < M; m++) {
        float v = warp_sum(__fadd_rn(acc[m], accm[m]));
        if (lane == 0) {
            if (bias) v = __fadd_rn(v, bias[n]);
            if (p.res) v = __fadd_rn(p.res[(size_t)m * p.ldr + n], v);
            y[(size_t)m * p.ldy + n] = v;
        }
    }
}

// One activation row, K <= 4096 (q|k|v, o, gate|up, lm_head of a 7B body): the lane's share of the activation row -- the same 4 x (32 int8,
// 2 group sums, d) for EVERY weight row -- lives in registers, and a persistent grid's warps walk rows w, w + W, ... with only the two
// 16-byte weight loads per super-block left on the load path.  ncu on k_gemv_q4k (profiles/r02_ncu_gemv_q4k.txt): DRAM traffic equals the
// algorithmic bytes, but the L1 data pipe is 70 % busy and the issue slots 65 % -- five of the seven loads per super-block and lane were
// activation gathers repeated for every row.  Same integers, same per-lane fp32 folding as k_gemv_q4k.
template <int DUMMY>
__global__ void __launch_bounds__(128, 4) k_gemv_q4k_areg(const KqGemvParams p) {
    const int lane = threadIdx.x & 31;
    const int W = gridDim.x * 4, w = blockIdx.x * 4 + (threadIdx.x >> 5);
    const int nb = p.nb, g = lane >> 3, t = lane & 7, j = t >> 1, h = t & 1;
    const int sh16 = 16 * (j & 1);
    const bool upper = j >= 2;
    if (w >= p.Ntot) return;
    auto load = [&](int n, uint4 (&hq)[4], uint4 (&wq)[4]) {
        const uint8_t *Wp = p.W[0];
        if (n >= p.N[0]) { n -= p.N[0]; Wp = p.W[1]; if (n >= p.N[1]) { n -= p.N[1]; Wp = p.W[2]; } }
        const uint8_t *row = Wp + (size_t)n * nb * Q4K_BYTES;
#pragma unroll
        for (int u = 0; u < 4; u++) {
            const uint8_t *blk = row + (size_t)min(4 * u + g, nb - 1) * Q4K_BYTES;
            hq[u] = ld_stream_u4(blk);
            wq[u] = ld_stream_u4(blk + 16 + t * 16);
        }
    };
    uint4 hq[4], wq[4];
    load(w, hq, wq);
    kq_pdl_launch();
    kq_pdl_wait(); // the activations are the predecessor's output; the first row's weights are already on their way
    int4 al[4], ah[4];
    int bl[4], bh[4];
    float dx[4];
#pragma unroll
    for (int u = 0; u < 4; u++) {
        const int i = min(4 * u + g, nb - 1);
        const uint8_t *xb = p.xq + (size_t)i * Q8K_PITCH;
        al[u] = *(const int4 *)(xb + 16 + (2 * j) * 32 + h * 16);
        ah[u] = *(const int4 *)(xb + 16 + (2 * j + 1) * 32 + h * 16);
        const int16_t *bs = (const int16_t *)(xb + 272);
        bl[u] = (int)bs[4 * j + h];
        bh[u] = (int)bs[4 * j + 2 + h];
        dx[u] = 4 * u + g < nb ? *(const float *)xb : 0.f; // a lane past the end of the row contributes with factor 0
    }
    auto row_ptr = [&](int n) {
        const uint8_t *Wp = p.W[0];
        if (n >= p.N[0]) { n -= p.N[0]; Wp = p.W[1]; if (n >= p.N[1]) { n -= p.N[1]; Wp = p.W[2]; } }
        return Wp + (size_t)n * nb * Q4K_BYTES;
    };
    for (int n = w; n < p.Ntot; n += W) {
        float acc = 0.f, accm = 0.f;
        const bool more = n + W < p.Ntot;
        const uint8_t *next = row_ptr(more ? n + W : n);
#pragma unroll
        for (int u = 0; u < 4; u++) {
            const uint32_t y2 = hq[u].y >> sh16, z2 = hq[u].z >> sh16, w2 = hq[u].w >> sh16; // get_scale_min_k4 for sub-blocks 2j, 2j+1 only
            const uint32_t sc2 = upper ? ((w2 & 0x0F0Fu) | ((y2 >> 2) & 0x3030u)) : (y2 & 0x3F3Fu);
            const uint32_t mn2 = upper ? (((w2 >> 4) & 0x0F0Fu) | ((z2 >> 2) & 0x3030u)) : (z2 & 0x3F3Fu);
            const int sc_lo = (int)(sc2 & 0xFFu), sc_hi = (int)(sc2 >> 8), mn_lo = (int)(mn2 & 0xFFu), mn_hi = (int)(mn2 >> 8);
            const float dw = __half2float(__ushort_as_half((unsigned short)(hq[u].x & 0xFFFFu)));
            const float dminw = __half2float(__ushort_as_half((unsigned short)(hq[u].x >> 16)));
            int is_lo = 0, is_hi = 0;
            is_lo = __dp4a((int)(wq[u].x & 0x0F0F0F0Fu), al[u].x, is_lo); is_hi = __dp4a((int)((wq[u].x >> 4) & 0x0F0F0F0Fu), ah[u].x, is_hi);
            is_lo = __dp4a((int)(wq[u].y & 0x0F0F0F0Fu), al[u].y, is_lo); is_hi = __dp4a((int)((wq[u].y >> 4) & 0x0F0F0F0Fu), ah[u].y, is_hi);
            is_lo = __dp4a((int)(wq[u].z & 0x0F0F0F0Fu), al[u].z, is_lo); is_hi = __dp4a((int)((wq[u].z >> 4) & 0x0F0F0F0Fu), ah[u].z, is_hi);
            is_lo = __dp4a((int)(wq[u].w & 0x0F0F0F0Fu), al[u].w, is_lo); is_hi = __dp4a((int)((wq[u].w >> 4) & 0x0F0F0F0Fu), ah[u].w, is_hi);
            if (more) { // this group's registers are free: the next row's same group starts its trip now
                const uint8_t *blk = next + (size_t)min(4 * u + g, nb - 1) * Q4K_BYTES;
                hq[u] = ld_stream_u4(blk);
                wq[u] = ld_stream_u4(blk + 16 + t * 16);
            }
            const int vi = sc_lo * is_lo + sc_hi * is_hi;
            const int vm = mn_lo * bl[u] + mn_hi * bh[u];
            acc = fmaf(__fmul_rn(dx[u], dw), (float)vi, acc);
            accm = fmaf(__fmul_rn(-dx[u], dminw), (float)vm, accm);
        }
        int nn = n, s = 0;
        if (nn >= p.N[0]) { nn -= p.N[0]; s = 1; if (nn >= p.N[1]) { nn -= p.N[1]; s = 2; } }
        const float *bias = s == 0 ? p.bias[0] : s == 1 ? p.bias[1] : p.bias[2];
        float *y = s == 0 ? p.y[0] : s == 1 ? p.y[1] : p.y[2];
        float v = warp_sum(__fadd_rn(acc, accm));
        if (lane == 0) {
            if (bias) v = __fadd_rn(v, bias[nn]);
            if (p.res) v = __fadd_rn(p.res[nn], v);
            y[nn] = v;
        }
    }
}

MLLMCU_API int mllmcu_gemv_q4k(int nW, const void *const *W, const int *N, const float *const *bias, float *const *y, int ldy, const void *xq_blocks,
                               int M, int K, const float *residual, int ldr, int flags, void *stream) {
    const int pdl = flags & 1, aligned = flags & 2, no_areg = flags & 4;
    if (aligned && ((uintptr_t)xq_blocks & 15)) return -1;
    if (nW < 1 || nW > 3 || M < 1 || M > 8 || K <= 0 || K % QK_K || ((uintptr_t)xq_blocks & 3)) return -1;
    if (residual && nW != 1) return -1;
    KqGemvParams p = {};
    p.nW = nW; p.nb = K / QK_K; p.ldy = ldy; p.ldr = ldr; p.xq = (const uint8_t *)xq_blocks; p.res = residual;
    for (int s = 0; s < 3; s++) p.N[s] = 0x7fffffff;
    long long tot = 0;
    for (int s = 0; s < nW; s++) {
        if (W[s] == nullptr || y[s] == nullptr || N[s] <= 0 || ((uintptr_t)W[s] & 15)) return -1;
        p.W[s] = (const uint8_t *)W[s]; p.bias[s] = bias ? bias[s] : nullptr; p.y[s] = y[s]; p.N[s] = N[s];
        tot += N[s];
    }
    if (tot > 0x7fffffff) return -1;
    p.Ntot = (int)tot;
    cudaLaunchConfig_t cfg;
    cudaLaunchAttribute attr[1];
    static const int areg_ctas = getenv("MLLMCU_Q4K_AREG_CTAS") ? atoi(getenv("MLLMCU_Q4K_AREG_CTAS")) : 4; // resident CTAs per SM the persistent grid is sized for; 0 = off
    if (M == 1 && aligned && p.nb <= 16 && !no_areg && areg_ctas > 0) { // one activation row held in registers
        kq_launch_cfg(cfg, attr, dim3((unsigned)std::min((p.Ntot + 3) / 4, cu_sm_count() * areg_ctas)), dim3(128), pdl, (cudaStream_t)stream);
        CU_TRY(cudaLaunchKernelEx(&cfg, k_gemv_q4k_areg<0>, p));
        return 0;
    }
    kq_launch_cfg(cfg, attr, dim3((unsigned)((p.Ntot + 3) / 4)), dim3(128), pdl, (cudaStream_t)stream);
    switch (M) {
#define CASE(m) case m: if (aligned) CU_TRY(cudaLaunchKernelEx(&cfg, k_gemv_q4k<m, true>, p)); else CU_TRY(cudaLaunchKernelEx(&cfg, k_gemv_q4k<m, false>, p)); break;
        CASE(1) CASE(2) CASE(3) CASE(4) CASE(5) CASE(6) CASE(7) CASE(8)
#undef CASE
    }
    return 0;
}

// ------------------------------------------------- exact integers per super-block (test / diagnostic hook)
// sumi[n][m][i] = sum_sb scale_sb * (q4 . q8), summ[n][m][i] = sum_sb min_sb * (bsums[2sb] + bsums[2sb+1])
__global__ void k_block_isums_q4_K(const uint8_t *__restrict__ W, const uint8_t *__restrict__ xq, int32_t *__restrict__ sumi, int32_t *__restrict__ summ,
                                   int M, int N, int nb) {
    const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= (long long)M * N * nb) return;
    const int i = (int)(idx % nb);
    const int m = (int)((idx / nb) % M);
    const int n = (int)(idx / ((long long)nb * M));
    const uint8_t *blk = W + ((size_t)n * nb + i) * Q4K_BYTES;
    const uint8_t *xb = xq + ((size_t)m * nb + i) * Q8K_BYTES;
    const int16_t *bs = (const int16_t *)(xb + 260);
    int si = 0, sm = 0;
    for (int c = 0; c < 8; c++) {
        int w[8], sc, mn, is = 0;
        kq_chunk_words<12>(blk, c, w);
        q4k_scale_min(blk + 4, c, sc, mn);
        const int *a = (const int *)(xb + 4 + c * 32);
#pragma unroll
        for (int j = 0; j < 8; j++) is = __dp4a(w[j], a[j], is);
        si += sc * is;
        sm += mn * ((int)bs[2 * c] + (int)bs[2 * c + 1]);
    }
    sumi[idx] = si;
    summ[idx] = sm;
}
MLLMCU_API int mllmcu_block_isums_q4_K_q8_K(const void *W, const void *xq_blocks, int32_t *sumi, int32_t *summ, int M, int N, int K, void *stream) {
    if (K % QK_K || M <= 0 || N <= 0 || ((uintptr_t)W & 15)) return -1;
    const int nb = K / QK_K;
    const long long tot = (long long)M * N * nb;
    k_block_isums_q4_K<<<(unsigned)((tot + 127) / 128), 128, 0, (cudaStream_t)stream>>>((const uint8_t *)W, (const uint8_t *)xq_blocks, sumi, summ, M, N, nb);
    CU_LAUNCH_CHECK();
    return 0;
}
