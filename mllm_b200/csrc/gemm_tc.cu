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
//   warps 5-20  : epilogue -- tcgen05.ld (32 lanes x TC_EPI_COLS = 32 columns per warp), scale, accumulate in registers,
//                 finally bias and store
// Pipelines: 4 smem stages of 2 Q blocks, 4 TMEM buffers; all hand-offs through mbarriers.
#include "common.cuh"
#include "quantize.cuh"
#include <stdlib.h>

#define TC_BM 128
#define TC_BN 128
#define TC_KB 2   // Q blocks per shared-memory stage
#define TC_NS 4   // shared-memory stages
#define TC_ND 4   // TMEM accumulator buffers (one Q block each)
#define TC_SR 16  // scale ring depth in Q blocks (> TC_NS*TC_KB + TC_ND: loaders never lap the epilogue)
#ifndef TC_EPI_WARPS
#define TC_EPI_WARPS 16 // 4 TMEM lane quarters x 4 column quarters of 32.  8 (x 64 columns) executes 30 % fewer instructions per Q block
                        // (the barrier wait, address arithmetic and loop cost ~56 of the 88 per warp whatever the width) and is SLOWER
                        // (98.9 against 86 us per 512 x 4096 x 4096, prefill 39.0 against 33.7 ms): the epilogue is bound by the latency
                        // of tcgen05.ld / LDS chains, which 16 warps hide better (profiles/r02_gemm_tc_epilogue_widths.txt)
#endif
#define TC_EPI_COLS (TC_BN / (TC_EPI_WARPS / 4)) // accumulator columns per epilogue warp
#define TC_THREADS ((5 + TC_EPI_WARPS) * 32)
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
// One lane of a CONVERGED warp (elect.sync).  The MMA-issuing warp runs its loop with all 32 lanes and lets the elected lane
// issue: inside an `if (lane == 0)` region the compiler has to wrap every uniform-operand instruction (UTCHMMA, UTCBAR) in an
// elect / R2UR.BROADCAST / BRA.U.ANY loop, which costs ~150 cycles per MMA against ~70 this way (profiles/micro/mma_rate.cu).
static __device__ __forceinline__ uint32_t tc_elect_one() {
    uint32_t pred;
    asm volatile("{\n.reg .b32 rx;\n.reg .pred px;\nelect.sync rx|px, 0xffffffff;\nselp.b32 %0, 1, 0, px;\n}" : "=r"(pred));
    return pred;
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
// Q8_0 values (bytes identical to quantize_row_q8_0) stored as fp16 integers, already in the shared-memory image the
// GEMM wants: for every 128-row tile, 16-byte chunks of 8 consecutive k ordered [k chunk][row] -- so the A operand of
// a pipeline stage (8 chunks x 128 rows) is ONE contiguous 16 KB range and moves with a single bulk copy.
//   qt[((mt * K/8 + kc) * 128 + r) * 8 + e]   row = mt*128 + r, k = kc*8 + e;   dt[(mt * nb + b) * 128 + r] = d_x
// Rows beyond `rows` in the last tile are written as zeros.
// Thread = 8 consecutive k of one row (one 16-byte chunk of the image); a warp = 8 rows x the 4 chunks of one Q block, so a row's
// 128 bytes are read by 4 adjacent lanes and each chunk column is written as 8 adjacent 16-byte stores (full sectors both ways).
__global__ void __launch_bounds__(256) k_quantize_q8_0_f16_tiled(const float *__restrict__ x, __half *__restrict__ qt, float *__restrict__ dt, int rows,
                                                                  int rows_pad, int nb, int ldx) {
    const int lane = threadIdx.x & 31, kc = lane & 3, rl = lane >> 2;
    const long long wid = (long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); // (row group of 8, block)
    const long long ngroups = (long long)(rows_pad >> 3) * nb;
    if (wid >= ngroups) return;
    const int b = (int)(wid % nb), row = (int)(wid / nb) * 8 + rl;
    const int mt = row >> 7, r = row & 127;
    float4 v0 = make_float4(0.f, 0.f, 0.f, 0.f), v1 = v0;
    if (row < rows) {
        const float *p = x + (size_t)row * ldx + (size_t)b * QK + kc * 8;
        v0 = *(const float4 *)p;
        v1 = *(const float4 *)(p + 4);
    }
    float amax = fmaxf(fmaxf(fmaxf(fabsf(v0.x), fabsf(v0.y)), fmaxf(fabsf(v0.z), fabsf(v0.w))), fmaxf(fmaxf(fabsf(v1.x), fabsf(v1.y)), fmaxf(fabsf(v1.z), fabsf(v1.w))));
    amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, 1));
    amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, 2));
    const __half2 h0 = __halves2half2(__int2half_rn(q8_0_quant(v0.x, amax)), __int2half_rn(q8_0_quant(v0.y, amax)));
    const __half2 h1 = __halves2half2(__int2half_rn(q8_0_quant(v0.z, amax)), __int2half_rn(q8_0_quant(v0.w, amax)));
    const __half2 h2 = __halves2half2(__int2half_rn(q8_0_quant(v1.x, amax)), __int2half_rn(q8_0_quant(v1.y, amax)));
    const __half2 h3 = __halves2half2(__int2half_rn(q8_0_quant(v1.z, amax)), __int2half_rn(q8_0_quant(v1.w, amax)));
    uint4 o;
    o.x = *(const uint32_t *)&h0; o.y = *(const uint32_t *)&h1; o.z = *(const uint32_t *)&h2; o.w = *(const uint32_t *)&h3;
    *(uint4 *)(qt + (((size_t)mt * nb * 4 + (size_t)b * 4 + kc) * 128 + r) * 8) = o;
    if (kc == 0) dt[((size_t)mt * nb + b) * 128 + r] = __half2float(q8_0_scale(amax));
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
#define TC_SMEM_WST (TC_SMEM_BARS + 256)          // raw weight staging: [2 groups][128 rows][144 bytes]
#define TC_SMEM_TOTAL (TC_SMEM_WST + 2 * 128 * 144)

__global__ void __launch_bounds__(TC_THREADS, 1)
k_gemm_q4_0_tc(const uint8_t *__restrict__ W, const __half *__restrict__ qt, const float *__restrict__ dt, const float *__restrict__ bias,
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
            for (int s = 0; s < TC_NS; s++) { tc_mbar_init(FULL(s), 129); tc_mbar_init(EMPTY(s), 1); } // 128 loader rows + the A bulk copy
            for (int s = 0; s < TC_ND; s++) { tc_mbar_init(DFULL(s), 1); tc_mbar_init(DFREE(s), TC_EPI_WARPS); }
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
        // ---------------- MMA issue: the whole warp runs the loop, the elected lane issues
        const uint32_t el = tc_elect_one();
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
                if (el) {
#pragma unroll
                    for (int i = 0; i < 2; i++) { // K = 16 halves = two 16-byte core-matrix columns
                        const uint32_t koff = (uint32_t)(blk * 4 + 2 * i) * (TC_BM * 16);
                        tc_mma_f16(tmem_base + (uint32_t)buf * TC_BN, tc_desc(a_base + koff, TC_BM * 16, 128),
                                   tc_desc(b_base + koff, TC_BN * 16, 128), TC_IDESC, (uint32_t)i);
                    }
                    tc_commit(DFULL(buf));
                }
                __syncwarp();
            }
            if (el) tc_commit(EMPTY(s)); // all MMAs that read this stage have completed when this arrives
            __syncwarp();
        }
    } else if (warp <= 4) {
        // ---------------- loaders: one weight row per thread; thread 0 also starts the A bulk copy of each stage
        const int r = tid - 32;
        const int n = n0 + r;
        const uint4 *wrow = (const uint4 *)(W + (size_t)n * rowb);
        const __half *qtile = qt + (size_t)blockIdx.y * nb * 4 * 128 * 8;
        const float *dtile = dt + (size_t)blockIdx.y * nb * 128;
        const int ngroups = nb / 8; // 8 Q blocks = 144 bytes of a weight row = 9 aligned 16-byte loads
        // raw weights go global -> shared with per-thread async copies one 8-block group ahead (each thread only ever
        // reads back its own row, so no cross-thread synchronisation is needed)
        uint8_t *wst = smem + TC_SMEM_WST + (size_t)r * 144;
        auto fetch = [&](int g) {
            const uint32_t dst = tc_smem_u32(wst + (size_t)(g & 1) * (128 * 144));
#pragma unroll
            for (int i = 0; i < 9; i++)
                asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst + 16u * i), "l"(wrow + (size_t)g * 9 + i) : "memory");
            asm volatile("cp.async.commit_group;" ::: "memory");
        };
        fetch(0);
        for (int g = 0; g < ngroups; g++) {
            if (g + 1 < ngroups) {
                fetch(g + 1);
                asm volatile("cp.async.wait_group 1;" ::: "memory");
            } else {
                asm volatile("cp.async.wait_group 0;" ::: "memory");
            }
            const uint32_t *cw = (const uint32_t *)(wst + (size_t)(g & 1) * (128 * 144));
            float sxc[8];
#pragma unroll
            for (int i = 0; i < 8; i++) sxc[i] = dtile[(size_t)(g * 8 + i) * 128 + r];
#pragma unroll
            for (int cc = 0; cc < 4; cc++) {
                const int c = g * 4 + cc, s = c % TC_NS, b0 = c * TC_KB;
                tc_mbar_wait(EMPTY(s), (uint32_t)(((c / TC_NS) & 1) ^ 1));
                uint8_t *sa = smem + (size_t)s * TC_STAGE_BYTES, *sb = sa + TC_OPER_BYTES;
                if (r == 0) { // activations of this stage: one contiguous 16 KB image
                    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(FULL(s)), "r"((uint32_t)TC_OPER_BYTES) : "memory");
                    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(tc_smem_u32(sa)),
                                 "l"(qtile + (size_t)c * (TC_KB * 4) * 128 * 8), "r"((uint32_t)TC_OPER_BYTES), "r"(FULL(s))
                                 : "memory");
                }
                const float sx0 = sxc[2 * cc], sx1 = sxc[2 * cc + 1];
                uint32_t w[9]; // 2 blocks = 36 bytes = 9 words
#pragma unroll
                for (int i = 0; i < 9; i++) w[i] = cw[9 * cc + i];
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
                dwS[(b0 % TC_SR) * 128 + r] = d0;
                dwS[((b0 + 1) % TC_SR) * 128 + r] = d1;
                dxS[(b0 % TC_SR) * 128 + r] = sx0;
                dxS[((b0 + 1) % TC_SR) * 128 + r] = sx1;
                asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); // generic-proxy stores -> visible to the tensor core
                tc_mbar_arrive(FULL(s));
            }
        }
    } else {
        // ---------------- epilogue: TMEM lane quarter = warp % 4, TC_EPI_COLS of the 128 columns per warp
        constexpr int NC = TC_EPI_COLS, NV = NC / 32;
        const int q = warp & 3, cq = (warp - 5) >> 2;
        const int r = 32 * q + lane, m = m0 + r;
        float2 acc[NC / 2]; // columns 2j, 2j+1 of this warp's NC
#pragma unroll
        for (int j = 0; j < NC / 2; j++) acc[j] = make_float2(0.f, 0.f);
        for (int b = 0; b < nb; b++) {
            const int buf = b % TC_ND;
            tc_mbar_wait(DFULL(buf), (uint32_t)((b / TC_ND) & 1));
            tc_fence_after();
            uint32_t v[NV][32];
#pragma unroll
            for (int h = 0; h < NV; h++) tc_ld32(tmem_base + ((uint32_t)(32 * q) << 16) + (uint32_t)(buf * TC_BN + cq * NC + h * 32), v[h]);
            const float sx = dxS[(b % TC_SR) * 128 + r];
            const float4 *dwp = (const float4 *)(dwS + (b % TC_SR) * 128 + cq * NC);
            tc_wait_ld();
            tc_fence_before();
            __syncwarp();
            if (lane == 0) tc_mbar_arrive(DFREE(buf)); // accumulator drained: the next block may overwrite it
            const float2 sx2 = make_float2(sx, sx);
            // packed fp32x2: one FMUL2 (d_w * d_x, rounded as the reference's product) + one FFMA2 per element pair
#pragma unroll
            for (int h = 0; h < NV; h++)
#pragma unroll
                for (int j4 = 0; j4 < 8; j4++) {
                    const float4 wv = dwp[h * 8 + j4];
                    float2 &a0 = acc[h * 16 + 2 * j4], &a1 = acc[h * 16 + 2 * j4 + 1];
                    a0 = __ffma2_rn(__fmul2_rn(make_float2(wv.x, wv.y), sx2), make_float2(__uint_as_float(v[h][4 * j4]), __uint_as_float(v[h][4 * j4 + 1])), a0);
                    a1 = __ffma2_rn(__fmul2_rn(make_float2(wv.z, wv.w), sx2), make_float2(__uint_as_float(v[h][4 * j4 + 2]), __uint_as_float(v[h][4 * j4 + 3])), a1);
                }
        }
        if (m < M) {
            const int nc = n0 + cq * NC;
            if (out_f16) {
                __half *yp = (__half *)y + (size_t)m * ldy + nc;
#pragma unroll
                for (int j = 0; j < NC / 2; j++) {
                    yp[2 * j] = __float2half_rn(bias ? acc[j].x + bias[nc + 2 * j] : acc[j].x);
                    yp[2 * j + 1] = __float2half_rn(bias ? acc[j].y + bias[nc + 2 * j + 1] : acc[j].y);
                }
            } else {
                float *yp = (float *)y + (size_t)m * ldy + nc;
#pragma unroll
                for (int j = 0; j < NC / 2; j++) {
                    yp[2 * j] = bias ? acc[j].x + bias[nc + 2 * j] : acc[j].x;
                    yp[2 * j + 1] = bias ? acc[j].y + bias[nc + 2 * j + 1] : acc[j].y;
                }
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 0) {
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512u) : "memory");
    }
}


// =====================================================================================================================
// Scaled-operand kernel (opt-in: MLLMCU_TC_EXACT=0; measured 94-99 us against 86 us for the kernel above on 512 x 4096 x 4096,
// see profiles/r01_gemm_tc_scaled.txt for what was tried and where its time goes).  The per-block rescale above costs one FMUL + one FFMA per output element and Q block
// on the CUDA cores (~2300 cycles per 128x128 tile and block against 128 cycles of tensor work).  This kernel moves the
// scales INTO the operands instead, without giving up exactness of the products:
//   a[m][k] = d_x[m][b] * q8[m][k]        (18 significant bits)  = a_hi + a_lo,  a_hi = fp16(a) (HMUL2),  a_lo = fma(d, q, -a_hi)  (exact)
//   w[n][k] = d_w[n][b] * (q4[n][k] - 8)  (14 significant bits)  = w_hi + w_lo,  likewise
//   y[m][n] = sum_k a_hi w_hi + a_lo w_hi + a_hi w_lo            (three fp16 MMAs per k-step, fp32 accumulation in TMEM)
// Every product is exact in the fp32 accumulator; the dropped a_lo*w_lo term is <= 2^-22 of its product (an fp32 rounding).
// q8 / d_x are the bit-exact Q8_0 quantisation of the row (the same split image the dp4a kernels read: int8 qs[M][K] +
// float d[M][K/32]), so the result is the reference's sum_b d_w d_x isum_b evaluated in a different fp32 order.
// The tensor core's fp32 accumulation truncates (measured: a bias of ~3e-8 of the sum per MMA step), so a chain is cut
// after TS_CB blocks (96 MMAs): the accumulator is drained to registers by the epilogue warps (IEEE adds) while the
// next chunk runs in the other TMEM buffer.
// Stage (one Q block, K = 32): weights [w_hi 8 KB][w_lo 8 KB] in shared memory, each the canonical K-major no-swizzle
// core-matrix image ([4 k-chunks][128 rows][16 B]); activations [a_hi | a_lo] in TENSOR MEMORY (32 columns: row = lane, two
// K elements per column), written by their loader threads with tcgen05.st and read by the MMA as its A operand.  Shared
// memory then carries only the weight stores and the B-operand reads (40 KB per block, under the 128 B/clk the SM has
// for the 384 cycles of MMA work; with both operands in shared memory that was 80 KB and the kernel was shared-memory bound:
// profiles/r01_ncu_gemm_tcs_v2.txt).  Both operands are unpacked by loader threads (one row each).
//   warps 0-15: workers (unpack weight + activation rows of "their" blocks, drain finished chains, write the result)
//   warp 16: TMEM alloc + MMA issue (6 tcgen05.mma per block).  Raw weight rows are staged 8 blocks at a time with cp.async
//   (arrival through cp.async.mbarrier.arrive); each worker prefetches its own int8 activation blocks into registers.
#define TS_NS 6  // pipeline stages (one Q block each): weights in shared memory, activations in TMEM
#define TS_CB 16 // Q blocks per accumulation chain
#define TS_STAGE_BYTES (2 * 4 * 128 * 16) // w_hi + w_lo: 16 KB
#define TS_THREADS (17 * 32)
#define TS_SMEM_BARS (TS_NS * TS_STAGE_BYTES)
#define TS_SMEM_STG (TS_SMEM_BARS + 512)
#define TS_NSLOT 4                          // staged groups of raw weight rows (8 Q blocks = 144 B per row) in flight
#define TS_STG_BYTES (128 * 144)
#define TS_SMEM_TOTAL (TS_SMEM_STG + TS_NSLOT * TS_STG_BYTES)
#define TS_TMEM_A 256 // first TMEM column of the activation ring (32 columns per stage); accumulators are columns 0-255

// 4 bytes holding values u -> (u - bias) * d as hi / lo half2 pairs; `off` = fp16 (1024 + bias) twice
static __device__ __forceinline__ void bytes4_scaled(uint32_t x, __half2 off, __half2 dd, uint32_t &h01, uint32_t &h23, uint32_t &l01, uint32_t &l23) {
    const uint32_t a = __byte_perm(x, 0x64646464u, 0x4140), b = __byte_perm(x, 0x64646464u, 0x4342); // 1024 + u, exact in fp16
    const __half2 va = __hsub2(*(const __half2 *)&a, off), vb = __hsub2(*(const __half2 *)&b, off);
    const __half2 ha = __hmul2(va, dd), hb = __hmul2(vb, dd);                                     // fp16(d * v)
    const __half2 la = __hfma2(va, dd, __hneg2(ha)), lb = __hfma2(vb, dd, __hneg2(hb));           // d * v - hi, exact
    h01 = *(const uint32_t *)&ha; h23 = *(const uint32_t *)&hb;
    l01 = *(const uint32_t *)&la; l23 = *(const uint32_t *)&lb;
}
// D[tmem] (+)= A[tmem] * B[smem]^T: the A operand is read from tensor memory (row = lane, two K elements per 32-bit column)
static __device__ __forceinline__ void tc_mma_f16_ts(uint32_t tmem_d, uint32_t tmem_a, uint64_t desc_b, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "setp.ne.b32 p, %4, 0;\n"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n"
        "}" ::"r"(tmem_d), "r"(tmem_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
        : "memory");
}
static __device__ __forceinline__ void tc_st32(uint32_t taddr, const uint32_t (&r)[32]) {
    asm volatile(
        "tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, %17, %18, %19, %20, "
        "%21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31, %32};" ::"r"(taddr),
        "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]), "r"(r[9]), "r"(r[10]), "r"(r[11]), "r"(r[12]),
        "r"(r[13]), "r"(r[14]), "r"(r[15]), "r"(r[16]), "r"(r[17]), "r"(r[18]), "r"(r[19]), "r"(r[20]), "r"(r[21]), "r"(r[22]), "r"(r[23]), "r"(r[24]),
        "r"(r[25]), "r"(r[26]), "r"(r[27]), "r"(r[28]), "r"(r[29]), "r"(r[30]), "r"(r[31])
        : "memory");
}

__global__ void __launch_bounds__(TS_THREADS, 1)
k_gemm_q4_0_tcs(const uint8_t *__restrict__ W, const int8_t *__restrict__ xq, const float *__restrict__ xd, const float *__restrict__ bias,
                void *__restrict__ y, int M, int N, int K, int ldy, int out_f16) {
    extern __shared__ __align__(1024) uint8_t smem[];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int nb = K / QK, rowb = nb * 18;
    const int n0 = blockIdx.x * TC_BN, m0 = blockIdx.y * TC_BM;
    uint64_t *bars = (uint64_t *)(smem + TS_SMEM_BARS);
    uint32_t *tmem_slot = (uint32_t *)(smem + TS_SMEM_BARS + 448);
    const uint32_t bar0 = tc_smem_u32(bars);
    auto FULL = [&](int s) { return bar0 + 8u * s; };
    auto EMPTY = [&](int s) { return bar0 + 8u * (TS_NS + s); };
    auto DFULL = [&](int s) { return bar0 + 8u * (2 * TS_NS + s); };
    auto DFREE = [&](int s) { return bar0 + 8u * (2 * TS_NS + 2 + s); };
    auto SFULL = [&](int s) { return bar0 + 8u * (2 * TS_NS + 4 + s); };
    auto SEMPTY = [&](int s) { return bar0 + 8u * (2 * TS_NS + 4 + TS_NSLOT + s); };

    if (warp == 16) {
        if (lane == 0) {
            for (int s = 0; s < TS_NS; s++) { tc_mbar_init(FULL(s), 128); tc_mbar_init(EMPTY(s), 1); }
            for (int s = 0; s < 2; s++) { tc_mbar_init(DFULL(s), 1); tc_mbar_init(DFREE(s), 16); }
            for (int s = 0; s < TS_NSLOT; s++) { tc_mbar_init(SFULL(s), 512); tc_mbar_init(SEMPTY(s), 16); }
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

    if (warp == 16) {
        // ---------------- MMA issue: the whole warp runs the loop, the elected lane issues
        const uint32_t el = tc_elect_one();
        for (int b = 0; b < nb; b++) {
            const int s = b % TS_NS, chunk = b / TS_CB, buf = chunk & 1, bi = b % TS_CB;
            if (bi == 0) { // this chain's accumulator must have been drained (two chunks ago)
                tc_mbar_wait(DFREE(buf), (uint32_t)(((chunk >> 1) & 1) ^ 1));
                tc_fence_after();
            }
            tc_mbar_wait(FULL(s), (uint32_t)((b / TS_NS) & 1));
            tc_fence_after();
            const uint32_t wh = tc_smem_u32(smem + (size_t)s * TS_STAGE_BYTES), wl = wh + 8192;
            const uint32_t a0 = tmem_base + TS_TMEM_A + (uint32_t)s * 32; // [hi k0 | hi k1 | lo k0 | lo k1], 8 columns each
            const uint32_t acc = tmem_base + (uint32_t)buf * TC_BN;
            if (el) {
#pragma unroll
                for (int i = 0; i < 2; i++) { // K = 16 halves: two 16-byte core-matrix columns of W, 8 TMEM columns of A
                    const uint32_t koff = (uint32_t)(2 * i) * (128 * 16);
                    tc_mma_f16_ts(acc, a0 + 8 * i, tc_desc(wh + koff, 128 * 16, 128), TC_IDESC, (uint32_t)((bi | i) != 0));
                    tc_mma_f16_ts(acc, a0 + 16 + 8 * i, tc_desc(wh + koff, 128 * 16, 128), TC_IDESC, 1u);
                    tc_mma_f16_ts(acc, a0 + 8 * i, tc_desc(wl + koff, 128 * 16, 128), TC_IDESC, 1u);
                }
                tc_commit(EMPTY(s));
                if (bi == TS_CB - 1 || b == nb - 1) tc_commit(DFULL(buf));
            }
            __syncwarp();
        }
    } else {
        // ---------------- 16 worker warps.  Warp w serves rows 32 (w % 4) .. + 31 (weight row AND activation row = TMEM lane) and the
        // Q blocks b with b % 4 == w / 4, so four warps complete a stage and the four groups work on four stages at once; all 16
        // drain a finished accumulation chain (32 columns each) and write the result.
        const int rq = warp & 3, bq = warp >> 2;
        const int r = 32 * rq + lane, m = m0 + r;
        const bool live = m < M;
        const int ngroups = nb / 8, nchunks = (nb + TS_CB - 1) / TS_CB;
        const int mc = live ? m : M - 1; // rows beyond M compute on a copy of the last row; their results are never stored
        const int8_t *asrc = xq + (size_t)mc * K;
        const float *dsrc = xd + (size_t)mc * nb;
        const __half2 off8 = __halves2half2(__ushort_as_half((unsigned short)0x6408), __ushort_as_half((unsigned short)0x6408));     // 1032
        const __half2 off128 = __halves2half2(__ushort_as_half((unsigned short)0x6480), __ushort_as_half((unsigned short)0x6480));   // 1152
        const uint32_t t_lane = tmem_base + ((uint32_t)(32 * rq) << 16);

        // raw weight rows of group g (8 Q blocks: 128 rows x 144 B) -> staging slot g % TS_NSLOT, 16-byte async copies dealt over
        // the 512 worker threads; each thread's arrival on SFULL fires when its own copies have landed
        auto stage = [&](int g) {
            const int slot = g % TS_NSLOT;
            tc_mbar_wait(SEMPTY(slot), (uint32_t)(((g / TS_NSLOT) & 1) ^ 1));
            const uint32_t dst = tc_smem_u32(smem + TS_SMEM_STG + (size_t)slot * TS_STG_BYTES);
            for (int c = tid; c < 128 * 9; c += 512) {
                const int row = c / 9, k = c - 9 * row;
                asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst + (uint32_t)(row * 144 + 16 * k)),
                             "l"(W + (size_t)(n0 + row) * rowb + (size_t)g * 144 + 16 * k) : "memory");
            }
            asm volatile("cp.async.mbarrier.arrive.noinc.shared::cta.b64 [%0];" ::"r"(SFULL(slot)) : "memory");
        };
        // this thread's own activation block of visit v (block 8 (v / 2) + bq + 4 (v % 2) of row r: 32 int8 + scale), fetched into
        // registers one visit (four block times) ahead
        uint4 na0, na1;
        float nd;
        auto fetch_a = [&](int v) {
            const int b = (v >> 1) * 8 + bq + 4 * (v & 1);
            na0 = ld_stream_u4(asrc + (size_t)b * 32);
            na1 = ld_stream_u4(asrc + (size_t)b * 32 + 16);
            nd = dsrc[b];
        };
        float acc[32]; // running sums of the drained chains (keeping them in shared memory instead was measured slower)
#pragma unroll
        for (int j = 0; j < 32; j++) acc[j] = 0.f;
        int drained = 0;
        auto drain = [&](int c) {
            const int buf = c & 1;
            tc_mbar_wait(DFULL(buf), (uint32_t)((c >> 1) & 1));
            tc_fence_after();
            uint32_t v[32];
            tc_ld32(t_lane + (uint32_t)(buf * TC_BN + bq * 32), v);
            tc_wait_ld();
#pragma unroll
            for (int j = 0; j < 32; j++) acc[j] = __fadd_rn(acc[j], __uint_as_float(v[j]));
            tc_fence_before();
            __syncwarp();
            if (lane == 0) tc_mbar_arrive(DFREE(buf));
        };

        for (int g = 0; g < TS_NSLOT - 1 && g < ngroups; g++) stage(g);
        fetch_a(0);
        for (int g = 0; g < ngroups; g++) {
            const int slot = g % TS_NSLOT;
            const uint8_t *base = smem + TS_SMEM_STG + (size_t)slot * TS_STG_BYTES;
            tc_mbar_wait(SFULL(slot), (uint32_t)((g / TS_NSLOT) & 1));
#pragma unroll
            for (int h = 0; h < 2; h++) {
                const int j = bq + 4 * h, b = g * 8 + j, s = b % TS_NS;
                const uint4 a0 = na0, a1 = na1;
                const float dcur = nd;
                if (2 * g + h + 1 < 2 * ngroups) fetch_a(2 * g + h + 1);
                while (drained < b / TS_CB - 1) drain(drained++); // chains two chunks back are complete: free their accumulator
                // ---- weights: 18 bytes at a 2-byte-aligned offset; read as aligned words
                uint32_t q[4];
                unsigned short dbits;
                {
                    const uint32_t *wp = (const uint32_t *)(base + (size_t)r * 144 + 18 * j - ((j & 1) ? 2 : 0));
                    const uint32_t w0 = wp[0], w1 = wp[1], w2 = wp[2], w3 = wp[3], w4 = wp[4];
                    if (j & 1) { dbits = (unsigned short)(w0 >> 16); q[0] = w1; q[1] = w2; q[2] = w3; q[3] = w4; }
                    else { dbits = (unsigned short)(w0 & 0xffffu); q[0] = __byte_perm(w0, w1, 0x5432); q[1] = __byte_perm(w1, w2, 0x5432); q[2] = __byte_perm(w2, w3, 0x5432); q[3] = __byte_perm(w3, w4, 0x5432); }
                }
                uint4 ch[4], cl[4]; // elements 0-7, 8-15 (low nibbles of bytes 0-7, 8-15), 16-23, 24-31 (high nibbles)
                {
                    const __half dh = __ushort_as_half(dbits);
                    const __half2 dd = __halves2half2(dh, dh);
                    uint32_t lo[4], hi[4];
#pragma unroll
                    for (int i = 0; i < 4; i++) { lo[i] = q[i] & 0x0F0F0F0Fu; hi[i] = (q[i] >> 4) & 0x0F0F0F0Fu; }
                    bytes4_scaled(lo[0], off8, dd, ch[0].x, ch[0].y, cl[0].x, cl[0].y); bytes4_scaled(lo[1], off8, dd, ch[0].z, ch[0].w, cl[0].z, cl[0].w);
                    bytes4_scaled(lo[2], off8, dd, ch[1].x, ch[1].y, cl[1].x, cl[1].y); bytes4_scaled(lo[3], off8, dd, ch[1].z, ch[1].w, cl[1].z, cl[1].w);
                    bytes4_scaled(hi[0], off8, dd, ch[2].x, ch[2].y, cl[2].x, cl[2].y); bytes4_scaled(hi[1], off8, dd, ch[2].z, ch[2].w, cl[2].z, cl[2].w);
                    bytes4_scaled(hi[2], off8, dd, ch[3].x, ch[3].y, cl[3].x, cl[3].y); bytes4_scaled(hi[3], off8, dd, ch[3].z, ch[3].w, cl[3].z, cl[3].w);
                }
                tc_mbar_wait(EMPTY(s), (uint32_t)(((b / TS_NS) & 1) ^ 1));
                tc_fence_after();
                {
                    uint8_t *sh = smem + (size_t)s * TS_STAGE_BYTES, *sl = sh + 8192;
#pragma unroll
                    for (int kc = 0; kc < 4; kc++) {
                        *(uint4 *)(sh + ((size_t)kc * 128 + r) * 16) = ch[kc];
                        *(uint4 *)(sl + ((size_t)kc * 128 + r) * 16) = cl[kc];
                    }
                    asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); // generic-proxy stores -> visible to the tensor core
                }
                // ---- activations: 32 int8 + scale -> [hi | lo] fp16 pairs in 32 TMEM columns of lane r
                {
                    const __half dh = __float2half_rn(dcur); // d_x came from an fp16: exact
                    const __half2 dd = __halves2half2(dh, dh);
                    const uint32_t wq[8] = {a0.x ^ 0x80808080u, a0.y ^ 0x80808080u, a0.z ^ 0x80808080u, a0.w ^ 0x80808080u,
                                            a1.x ^ 0x80808080u, a1.y ^ 0x80808080u, a1.z ^ 0x80808080u, a1.w ^ 0x80808080u}; // q + 128 as unsigned bytes
                    uint32_t v[32]; // columns 0-15: hi (two elements each, k ascending), 16-31: lo
#pragma unroll
                    for (int i = 0; i < 8; i++) bytes4_scaled(wq[i], off128, dd, v[2 * i], v[2 * i + 1], v[16 + 2 * i], v[16 + 2 * i + 1]);
                    tc_st32(t_lane + TS_TMEM_A + (uint32_t)s * 32, v);
                    asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
                    tc_fence_before();
                }
                tc_mbar_arrive(FULL(s));
                // refill the slot group g - 1 used (every warp left it at least four blocks ago; unused so far when g == 0) with group g + TS_NSLOT - 1
                if (h == 0 && g + TS_NSLOT - 1 < ngroups) stage(g + TS_NSLOT - 1);
            }
            __syncwarp();
            if (lane == 0) tc_mbar_arrive(SEMPTY(slot)); // this warp no longer reads the staged group
        }
        while (drained < nchunks) drain(drained++);
        if (live) {
            const int nc = n0 + bq * 32;
            if (out_f16) {
                __half *yp = (__half *)y + (size_t)m * ldy + nc;
#pragma unroll
                for (int j = 0; j < 32; j++) yp[j] = __float2half_rn(bias ? acc[j] + bias[nc + j] : acc[j]);
            } else {
                float *yp = (float *)y + (size_t)m * ldy + nc;
#pragma unroll
                for (int j = 0; j < 32; j++) yp[j] = bias ? acc[j] + bias[nc + j] : acc[j];
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 16) {
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512u) : "memory");
    }
}

// scratch: tile-major fp16 activations (rows padded to 128), two planes for the scaled-operand kernel
size_t gemm_tc_workspace_bytes(int M, int K) {
    const size_t mp = ((size_t)M + TC_BM - 1) / TC_BM * TC_BM;
    return (mp * K * 2 + 255) / 256 * 256 + (mp * (K / QK) * 4 + 255) / 256 * 256;
}
bool gemm_tc_supported(int wtype, int M, int N, int K) {
    static int off = -1;
    if (off < 0) { const char *e = getenv("MLLMCU_NO_TC_GEMM"); off = (e && e[0] == '1') ? 1 : 0; }
    // K % 256: weight rows are read in 144-byte (8-block) groups that must start 16-byte aligned
    return !off && wtype == 2 && M > 8 && N % TC_BN == 0 && K % 256 == 0;
}
static bool tc_exact() { // MLLMCU_TC_EXACT=0 selects the scaled-operand kernel below (read per call: tests compare both)
    const char *ex = getenv("MLLMCU_TC_EXACT");
    return !(ex && ex[0] == '0');
}
// which activation image the workspace holds after a call: 3 = split Q8_0 (scaled-operand kernel; same as the dp4a kernels),
// 2 = tile-major integer fp16 + scales (exact kernel)
int gemm_tc_quant_kind() { return tc_exact() ? 2 : 3; }
// skip_quant: the workspace already holds this kernel's image of the same activations (an earlier Linear on the same input)
int launch_gemm_q4_0_tc(const void *W, const float *x, int ldx, const float *bias, void *y, int M, int N, int K, int ldy, int out_f16,
                        void *ws, int skip_quant, cudaStream_t st) {
    const int mp = (M + TC_BM - 1) / TC_BM * TC_BM;
    __half *qt = (__half *)ws;
    if (!tc_exact()) {
        // activations: the split Q8_0 image shared with the dp4a kernels (int8 qs[M][K] at ws, float d[M][K/32] after it)
        int8_t *xq = (int8_t *)ws;
        float *xd = (float *)((char *)ws + ((size_t)M * K + 255) / 256 * 256);
        if (!skip_quant) {
            const int rc = launch_quantize_q8_0_split(x, xq, xd, M, K, ldx, st);
            if (rc) return rc;
        }
        static bool configured_s = false;
        if (!configured_s) {
            CU_TRY(cudaFuncSetAttribute(k_gemm_q4_0_tcs, cudaFuncAttributeMaxDynamicSharedMemorySize, TS_SMEM_TOTAL));
            configured_s = true;
        }
        const dim3 grid(N / TC_BN, mp / TC_BM);
        k_gemm_q4_0_tcs<<<grid, TS_THREADS, TS_SMEM_TOTAL, st>>>((const uint8_t *)W, xq, xd, bias, y, M, N, K, ldy, out_f16);
        CU_LAUNCH_CHECK();
        return 0;
    }
    float *dt = (float *)((char *)ws + ((size_t)mp * K * 2 + 255) / 256 * 256);
    const int nb = K / QK;
    if (!skip_quant) {
        const long long groups = (long long)(mp >> 3) * nb; // a warp per 8 rows x 1 block
        k_quantize_q8_0_f16_tiled<<<(unsigned)((groups + 7) / 8), 256, 0, st>>>(x, qt, dt, M, mp, nb, ldx);
        CU_LAUNCH_CHECK();
    }
    static bool configured = false;
    if (!configured) {
        CU_TRY(cudaFuncSetAttribute(k_gemm_q4_0_tc, cudaFuncAttributeMaxDynamicSharedMemorySize, TC_SMEM_TOTAL));
        configured = true;
    }
    const dim3 grid(N / TC_BN, mp / TC_BM);
    k_gemm_q4_0_tc<<<grid, TC_THREADS, TC_SMEM_TOTAL, st>>>((const uint8_t *)W, qt, dt, bias, y, M, N, K, ldy, out_f16);
    CU_LAUNCH_CHECK();
    return 0;
}
