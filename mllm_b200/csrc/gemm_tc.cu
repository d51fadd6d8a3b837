// gemm_tc.cu -- prefill Linear (M > 8 rows) on the 5th-generation tensor cores (tcgen05 + TMEM).
//
// Reference arithmetic: mat_mul generic path (compute/Matmul.cpp:77-120,219-276):
//   y[m][n] = sum_b fma(d_w[n][b] * d_x[m][b], (float)isum[m][n][b], acc)  (+ bias),
//   isum = sum_{k in block b} (q4[n][k] - 8) * q8[m][k]   (vec_dot_q4_0_q8_0, ggml/VecDotQ4.cpp:514-544)
// with q8 / d_x the Q8_0 quantisation of the activation row (quantize_row_q8_0, ggml/QuantizeQ8.cpp:106-160).
//
// The scale changes every 32 elements of K, so the tensor core can only produce PER-BLOCK integer sums; the
// scale-and-accumulate runs on the CUDA cores.  To make that epilogue as cheap as possible the operands are fed
// as fp16 holding the exact integers (q4-8 in [-8,7], q8 in [-127,127]) with fp32 accumulation: every product
// and every partial sum (< 2^15) is exact, so the accumulator IS float(isum) and no int->float conversion is
// needed -- one FMUL (d_w*d_x) and one FFMA per element and block, same operations as the reference.
//
// CTA = 128 activation rows x 128 weight rows, one tcgen05.mma pair (K = 2 x 16) per Q block into one of four
// TMEM accumulator buffers (4 x 128 columns = the whole 512-column TMEM):
//   warp 0      : TMEM alloc, single-thread MMA issue, tcgen05.commit -> mbarriers
//   warps 1-4   : loaders -- Q4_0 rows -> exact fp16 (q-8) into the canonical no-swizzle K-major core-matrix
//                 layout (generic-proxy stores + fence.proxy.async), fp16 activation rows, d_w / d_x rings
//   warps 5-12  : epilogue -- tcgen05.ld (32 lanes x 64 columns per warp), scale, accumulate in registers,
//                 finally bias and store
// Pipelines: 4 smem stages of 2 Q blocks, 4 TMEM buffers; all hand-offs through mbarriers.
#include "common.cuh"
#include <stdlib.h>

#define TC_BM 128
#define TC_BN 128
#define TC_KB 2   // Q blocks per shared-memory stage
#define TC_NS 4   // shared-memory stages
#define TC_ND 4   // TMEM accumulator buffers (one Q block each)
#define TC_SR 16  // scale ring depth in Q blocks (> TC_NS*TC_KB + TC_ND: loaders never lap the epilogue)
#define TC_THREADS (13 * 32)
#define TC_STAGE_BYTES (2 * TC_KB * 4 * TC_BM * 16) // A + B: KB blocks x 4 16-byte K chunks x 128 rows
#define TC_OPER_BYTES (TC_KB * 4 * TC_BM * 16)

static __device__ __forceinline__ uint32_t tc_smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
static __device__ __forceinline__ void tc_mbar_init(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
static __device__ __forceinline__ void tc_mbar_arrive(uint32_t bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
// bounded spin: a protocol error traps instead of hanging the device
static __device__ __forceinline__ void tc_mbar_wait(uint32_t bar, uint32_t parity) {
    for (int spin = 0;; spin++) {
        uint32_t ok;
        asm volatile(
            "{\n"
            ".reg .pred P1;\n"
            "mbarrier.try_wait.parity.shared::cta.b64 P1, [%1], %2;\n"
            "selp.u32 %0, 1, 0, P1;\n"
            "}"
            : "=r"(ok)
            : "r"(bar), "r"(parity)
            : "memory");
        if (ok) return;
        if (spin > (1 << 26)) __trap();
    }
}
static __device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
static __device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
static __device__ __forceinline__ void tc_commit(uint32_t bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
// D[tmem] (+)= A[smem] * B[smem]^T, fp16 operands, fp32 accumulate, M=128, N=128, K=16
static __device__ __forceinline__ void tc_mma_f16(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "setp.ne.b32 p, %4, 0;\n"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n"
        "}" ::"r"(tmem_d), "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
        : "memory");
}
// shared-memory matrix descriptor, K-major, no swizzle: core matrix = 8 rows x 16 bytes stored contiguously;
// LBO = byte distance between the two K-adjacent core matrices, SBO = byte distance between 8-row groups
// (cute/atom/mma_traits_sm100.hpp make_umma_desc<Major::K>, LayoutType::INTERLEAVE)
static __device__ __forceinline__ uint64_t tc_desc(uint32_t saddr, uint32_t lbo, uint32_t sbo) {
    return (uint64_t)((saddr >> 4) & 0x3FFF) | ((uint64_t)((lbo >> 4) & 0x3FFF) << 16) | ((uint64_t)((sbo >> 4) & 0x3FFF) << 32) | (1ull << 46);
}
// instruction descriptor (cute/arch/mma_sm100_desc.hpp InstrDescriptor): D fp32, A/B fp16, both K-major
#define TC_IDESC ((1u << 4) | ((uint32_t)(TC_BN >> 3) << 17) | ((uint32_t)(TC_BM >> 4) << 24))

// 32 lanes x 32 consecutive 32-bit columns -> 32 registers per thread
static __device__ __forceinline__ void tc_ld32(uint32_t taddr, uint32_t (&r)[32]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, %17, %18, %19, "
        "%20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]), "=r"(r[9]), "=r"(r[10]),
          "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]),
          "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]),
          "=r"(r[31])
        : "r"(taddr));
}
static __device__ __forceinline__ void tc_wait_ld() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

// ------------------------------------------------------------------ activation quantisation for this kernel
// Q8_0 values (bytes identical to quantize_row_q8_0) stored as fp16 integers qh[rows][k], scales d[rows][nb] fp32.
__global__ void k_quantize_q8_0_f16(const float *__restrict__ x, __half *__restrict__ qh, float *__restrict__ d, int rows, int nb, int ldx) {
    const int lane = threadIdx.x & 31;
    const long long wid = (long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (wid >= (long long)rows * nb) return;
    const int r = (int)(wid / nb), b = (int)(wid % nb);
    const float v = x[(size_t)r * ldx + (size_t)b * QK + lane];
    const float amax = warp_max(fabsf(v));
    qh[((size_t)r * nb + b) * QK + lane] = __int2half_rn(q8_0_quant(v, amax));
    if (lane == 0) d[(size_t)r * nb + b] = __half2float(q8_0_scale(amax));
}

// 4 bytes holding values 0..15 -> two half2 (value - 8), exact: 0x6400 | v is the fp16 1024 + v
static __device__ __forceinline__ void nib4_to_h4(uint32_t x, uint32_t &h01, uint32_t &h23) {
    const uint32_t a = __byte_perm(x, 0x64646464u, 0x4140), b = __byte_perm(x, 0x64646464u, 0x4342);
    const __half2 off = __halves2half2(__ushort_as_half((unsigned short)0x6408), __ushort_as_half((unsigned short)0x6408)); // 1032
    const __half2 ra = __hsub2(*(const __half2 *)&a, off), rb = __hsub2(*(const __half2 *)&b, off);
    h01 = *(const uint32_t *)&ra;
    h23 = *(const uint32_t *)&rb;
}

#define TC_SMEM_SCALES (TC_NS * TC_STAGE_BYTES)
#define TC_SMEM_BARS (TC_SMEM_SCALES + 2 * TC_SR * 128 * 4)
#define TC_SMEM_TOTAL (TC_SMEM_BARS + 256)

__global__ void __launch_bounds__(TC_THREADS, 1)
k_gemm_q4_0_tc(const uint8_t *__restrict__ W, const __half *__restrict__ qh, const float *__restrict__ dx, const float *__restrict__ bias,
               void *__restrict__ y, int M, int N, int K, int ldy, int out_f16) {
    extern __shared__ __align__(1024) uint8_t smem[];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int nb = K / QK, nchunks = nb / TC_KB, rowb = nb * 18;
    const int n0 = blockIdx.x * TC_BN, m0 = blockIdx.y * TC_BM;
    float *dwS = (float *)(smem + TC_SMEM_SCALES);
    float *dxS = dwS + TC_SR * 128;
    uint64_t *bars = (uint64_t *)(smem + TC_SMEM_BARS);
    uint32_t *tmem_slot = (uint32_t *)(smem + TC_SMEM_BARS + 128);
    const uint32_t bar0 = tc_smem_u32(bars);
    auto FULL = [&](int s) { return bar0 + 8u * s; };
    auto EMPTY = [&](int s) { return bar0 + 8u * (TC_NS + s); };
    auto DFULL = [&](int s) { return bar0 + 8u * (2 * TC_NS + s); };
    auto DFREE = [&](int s) { return bar0 + 8u * (2 * TC_NS + TC_ND + s); };

    if (warp == 0) {
        if (lane == 0) {
            for (int s = 0; s < TC_NS; s++) { tc_mbar_init(FULL(s), 128); tc_mbar_init(EMPTY(s), 1); }
            for (int s = 0; s < TC_ND; s++) { tc_mbar_init(DFULL(s), 1); tc_mbar_init(DFREE(s), 8); }
            asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        }
        __syncwarp();
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tc_smem_u32(tmem_slot)), "r"(512u) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0) {
        // ---------------- MMA issue (one thread)
        if (lane == 0) {
            for (int c = 0; c < nchunks; c++) {
                const int s = c % TC_NS;
                tc_mbar_wait(FULL(s), (uint32_t)((c / TC_NS) & 1));
                tc_fence_after();
                const uint32_t a_base = tc_smem_u32(smem + (size_t)s * TC_STAGE_BYTES), b_base = a_base + TC_OPER_BYTES;
#pragma unroll
                for (int blk = 0; blk < TC_KB; blk++) {
                    const int b = c * TC_KB + blk, buf = b % TC_ND;
                    tc_mbar_wait(DFREE(buf), (uint32_t)(((b / TC_ND) & 1) ^ 1));
                    tc_fence_after();
#pragma unroll
                    for (int i = 0; i < 2; i++) { // K = 16 halves = two 16-byte core-matrix columns
                        const uint32_t koff = (uint32_t)(blk * 4 + 2 * i) * (TC_BM * 16);
                        tc_mma_f16(tmem_base + (uint32_t)buf * TC_BN, tc_desc(a_base + koff, TC_BM * 16, 128),
                                   tc_desc(b_base + koff, TC_BN * 16, 128), TC_IDESC, (uint32_t)i);
                    }
                    tc_commit(DFULL(buf));
                }
                tc_commit(EMPTY(s)); // all MMAs that read this stage have completed when this arrives
            }
        }
    } else if (warp <= 4) {
        // ---------------- loaders: one operand row per thread
        const int r = tid - 32;
        const int n = n0 + r, m = m0 + r;
        const uint8_t *wrow = W + (size_t)n * rowb;
        for (int c = 0; c < nchunks; c++) {
            const int s = c % TC_NS;
            tc_mbar_wait(EMPTY(s), (uint32_t)(((c / TC_NS) & 1) ^ 1));
            uint8_t *sa = smem + (size_t)s * TC_STAGE_BYTES, *sb = sa + TC_OPER_BYTES;
            // weights: 2 blocks = 36 bytes = 9 words (4-byte aligned because c * 36 and the row length are)
            const uint32_t *wp = (const uint32_t *)(wrow + (size_t)c * (TC_KB * 18));
            uint32_t w[9];
#pragma unroll
            for (int i = 0; i < 9; i++) w[i] = __ldg(wp + i);
            uint32_t q[2][4];
#pragma unroll
            for (int j = 0; j < 4; j++) { q[0][j] = __byte_perm(w[j], w[j + 1], 0x5432); q[1][j] = w[5 + j]; }
            const float d0 = __half2float(__ushort_as_half((unsigned short)(w[0] & 0xffffu)));
            const float d1 = __half2float(__ushort_as_half((unsigned short)(w[4] >> 16)));
#pragma unroll
            for (int blk = 0; blk < 2; blk++) {
                uint32_t lo[4], hi[4];
#pragma unroll
                for (int j = 0; j < 4; j++) { lo[j] = q[blk][j] & 0x0F0F0F0Fu; hi[j] = (q[blk][j] >> 4) & 0x0F0F0F0Fu; }
                uint4 ch[4]; // elements 0-7, 8-15 (low nibbles of bytes 0-7, 8-15), 16-23, 24-31 (high nibbles)
                nib4_to_h4(lo[0], ch[0].x, ch[0].y); nib4_to_h4(lo[1], ch[0].z, ch[0].w);
                nib4_to_h4(lo[2], ch[1].x, ch[1].y); nib4_to_h4(lo[3], ch[1].z, ch[1].w);
                nib4_to_h4(hi[0], ch[2].x, ch[2].y); nib4_to_h4(hi[1], ch[2].z, ch[2].w);
                nib4_to_h4(hi[2], ch[3].x, ch[3].y); nib4_to_h4(hi[3], ch[3].z, ch[3].w);
#pragma unroll
                for (int kc = 0; kc < 4; kc++) *(uint4 *)(sb + ((size_t)(blk * 4 + kc) * TC_BN + r) * 16) = ch[kc];
            }
            const int b0 = c * TC_KB;
            dwS[(b0 % TC_SR) * 128 + r] = d0;
            dwS[((b0 + 1) % TC_SR) * 128 + r] = d1;
            // activations: 64 fp16 integers = 8 chunks of 16 bytes
            if (m < M) {
                const uint4 *ap = (const uint4 *)(qh + (size_t)m * K + (size_t)c * (TC_KB * QK));
#pragma unroll
                for (int j = 0; j < 8; j++) *(uint4 *)(sa + ((size_t)j * TC_BM + r) * 16) = __ldg(ap + j);
                dxS[(b0 % TC_SR) * 128 + r] = dx[(size_t)m * nb + b0];
                dxS[((b0 + 1) % TC_SR) * 128 + r] = dx[(size_t)m * nb + b0 + 1];
            } else {
                const uint4 z = make_uint4(0, 0, 0, 0);
#pragma unroll
                for (int j = 0; j < 8; j++) *(uint4 *)(sa + ((size_t)j * TC_BM + r) * 16) = z;
                dxS[(b0 % TC_SR) * 128 + r] = 0.f;
                dxS[((b0 + 1) % TC_SR) * 128 + r] = 0.f;
            }
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); // generic-proxy stores -> visible to the tensor core
            tc_mbar_arrive(FULL(s));
        }
    } else {
        // ---------------- epilogue: TMEM lane quarter = warp % 4, 64 of the 128 columns per warp
        const int q = warp & 3, hc = (warp - 5) >> 2;
        const int r = 32 * q + lane, m = m0 + r;
        float acc[64];
#pragma unroll
        for (int j = 0; j < 64; j++) acc[j] = 0.f;
        for (int b = 0; b < nb; b++) {
            const int buf = b % TC_ND;
            tc_mbar_wait(DFULL(buf), (uint32_t)((b / TC_ND) & 1));
            tc_fence_after();
            const uint32_t taddr = tmem_base + ((uint32_t)(32 * q) << 16) + (uint32_t)(buf * TC_BN + hc * 64);
            uint32_t v0[32], v1[32];
            tc_ld32(taddr, v0);
            tc_ld32(taddr + 32, v1);
            tc_wait_ld();
            tc_fence_before();
            __syncwarp();
            if (lane == 0) tc_mbar_arrive(DFREE(buf)); // accumulator drained: the next block may overwrite it
            const float sx = dxS[(b % TC_SR) * 128 + r];
            const float4 *dwp = (const float4 *)(dwS + (b % TC_SR) * 128 + hc * 64);
#pragma unroll
            for (int j4 = 0; j4 < 8; j4++) {
                const float4 wv = dwp[j4];
                acc[4 * j4 + 0] = fmaf(__fmul_rn(wv.x, sx), __uint_as_float(v0[4 * j4 + 0]), acc[4 * j4 + 0]);
                acc[4 * j4 + 1] = fmaf(__fmul_rn(wv.y, sx), __uint_as_float(v0[4 * j4 + 1]), acc[4 * j4 + 1]);
                acc[4 * j4 + 2] = fmaf(__fmul_rn(wv.z, sx), __uint_as_float(v0[4 * j4 + 2]), acc[4 * j4 + 2]);
                acc[4 * j4 + 3] = fmaf(__fmul_rn(wv.w, sx), __uint_as_float(v0[4 * j4 + 3]), acc[4 * j4 + 3]);
            }
#pragma unroll
            for (int j4 = 0; j4 < 8; j4++) {
                const float4 wv = dwp[8 + j4];
                acc[32 + 4 * j4 + 0] = fmaf(__fmul_rn(wv.x, sx), __uint_as_float(v1[4 * j4 + 0]), acc[32 + 4 * j4 + 0]);
                acc[32 + 4 * j4 + 1] = fmaf(__fmul_rn(wv.y, sx), __uint_as_float(v1[4 * j4 + 1]), acc[32 + 4 * j4 + 1]);
                acc[32 + 4 * j4 + 2] = fmaf(__fmul_rn(wv.z, sx), __uint_as_float(v1[4 * j4 + 2]), acc[32 + 4 * j4 + 2]);
                acc[32 + 4 * j4 + 3] = fmaf(__fmul_rn(wv.w, sx), __uint_as_float(v1[4 * j4 + 3]), acc[32 + 4 * j4 + 3]);
            }
        }
        if (m < M) {
            const int nc = n0 + hc * 64;
            if (out_f16) {
                __half *yp = (__half *)y + (size_t)m * ldy + nc;
#pragma unroll
                for (int j = 0; j < 64; j++) yp[j] = __float2half_rn(bias ? acc[j] + bias[nc + j] : acc[j]);
            } else {
                float *yp = (float *)y + (size_t)m * ldy + nc;
#pragma unroll
                for (int j = 0; j < 64; j++) yp[j] = bias ? acc[j] + bias[nc + j] : acc[j];
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 0) {
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512u) : "memory");
    }
}

// qh/dx scratch: rows x K fp16 + rows x K/32 fp32
size_t gemm_tc_workspace_bytes(int M, int K) {
    return ((size_t)M * K * 2 + 255) / 256 * 256 + ((size_t)M * (K / QK) * 4 + 255) / 256 * 256;
}
bool gemm_tc_supported(int wtype, int M, int N, int K) {
    static int off = -1;
    if (off < 0) { const char *e = getenv("MLLMCU_NO_TC_GEMM"); off = (e && e[0] == '1') ? 1 : 0; }
    return !off && wtype == 2 && M > 8 && N % TC_BN == 0 && K % (TC_KB * QK) == 0;
}
int launch_gemm_q4_0_tc(const void *W, const float *x, int ldx, const float *bias, void *y, int M, int N, int K, int ldy, int out_f16,
                        void *ws, cudaStream_t st) {
    __half *qh = (__half *)ws;
    float *dx = (float *)((char *)ws + ((size_t)M * K * 2 + 255) / 256 * 256);
    const int nb = K / QK;
    const long long warps = (long long)M * nb;
    k_quantize_q8_0_f16<<<(unsigned)((warps + 7) / 8), 256, 0, st>>>(x, qh, dx, M, nb, ldx);
    CU_LAUNCH_CHECK();
    static bool configured = false;
    if (!configured) {
        CU_TRY(cudaFuncSetAttribute(k_gemm_q4_0_tc, cudaFuncAttributeMaxDynamicSharedMemorySize, TC_SMEM_TOTAL));
        configured = true;
    }
    const dim3 grid(N / TC_BN, (M + TC_BM - 1) / TC_BM);
    k_gemm_q4_0_tc<<<grid, TC_THREADS, TC_SMEM_TOTAL, st>>>((const uint8_t *)W, qh, dx, bias, y, M, N, K, ldy, out_f16);
    CU_LAUNCH_CHECK();
    return 0;
}
