// gemm_mma_q4k.cu -- Q4_K Linear with 9+ activation rows (prefill, batched decode) on the tensor cores.
//
// Reference arithmetic: mat_mul's generic path (compute/Matmul.cpp:104-115,223-276) with quantize_row_q8_K (ggml/QuantizeQ8.cpp:
// 216-266) and vec_dot_q4_K_q8_K (ggml/VecDotQ4.cpp:32-345): per 256-element super-block TWO exact integers,
//   sumi = sum_{j<8} scale_j * (q4_j . q8_j)            (q4 in 0..15, q8 in -128..127, scale_j 6 bits)
//   summ = sum_{j<8} min_j * (bsums[2j] + bsums[2j+1])  (the group sums quantize_row_q8_K stores)
// and y += (d_x * fp16(d_w)) * sumi - (d_x * fp16(dmin_w)) * summ in fp32.  The CUDA-core kernels of kquants.cu form the same
// integers with dp4a (k_gemm_kq: "20 Tflop/s-equivalent"); here the 32-element dots are ONE mma.sync.m16n8k32.u8.s8 per 16 weight rows
// x 8 activation rows x sub-block (weights as the A operand: nibbles as stored, no offset to remove -- the min term carries it), its
// int32 results are multiplied by the 6-bit scales and summed with integer multiply-adds, exactly the reference's integers, and only the
// per-super-block FMA pair runs in fp32.  Activations are the NATIVE block_q8_K rows kquants.cu's quantiser writes (bytes identical to
// the reference's): the B fragments, the group sums and d_x are read from them directly.
//
// Layout: a tile = 16 consecutive weight rows; a stage = KC super-blocks of those rows (one cp.async.bulk per row into a padded pitch,
// so the eight rows of a fragment start in different banks), staged through a ring filled ahead of the dependency wait.  16 warps =
// MG groups of MTW x 8 activation rows x KS K-splits over the super-blocks of a stage; the partial sums of the K-splits meet in shared
// memory in a fixed order.  M rows beyond 64 (128 for short K) are served by further launches over the same weights (they stay in L2).
#include "common.cuh"
#include "gemv_common.cuh"
#include "kq_common.cuh"
#include "mllmcu.h"
#include <stdlib.h>

#define KQ_WARPS 16
#define KQ_Q4K 144

struct GemmQ4KParams {
    const uint8_t *W;   // [N][nsb] block_q4_K
    const uint8_t *xq;  // fragment-major activation image (kquants.cu, k_quantize_q8_K_frag): QS region
    const uint8_t *xdx, *xbs; // its DX and BS regions
    int grp0, G;        // first 8-row group of this launch, groups in the image
    const float *bias;
    void *y;
    int N, M, nsb, ldy, out_f16;
    int KC, nchunk, stages, pitch; // super-blocks per stage, stages per tile, ring depth, bytes between rows of a stage
};

static __device__ __forceinline__ void mma_u8s8(int (&c)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t b0, uint32_t b1) {
    asm volatile("mma.sync.aligned.m16n8k32.row.col.s32.u8.s8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%10,%11,%12,%13};"
                 : "=r"(c[0]), "=r"(c[1]), "=r"(c[2]), "=r"(c[3])
                 : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1), "r"(0), "r"(0), "r"(0), "r"(0));
}

template <int MTW, int MG, int KS>
__global__ void __launch_bounds__(KQ_WARPS * 32) k_gemm_mma_q4k(const GemmQ4KParams p) {
    static_assert(MG * KS == KQ_WARPS, "16 warps");
    constexpr int Mp = MG * MTW * 8;
    extern __shared__ __align__(128) uint8_t smem[];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int g = lane >> 2, t = lane & 3;
    const int ks = warp % KS, mg = warp / KS;
    const int nsb = p.nsb, S = p.stages, KC = p.KC, nchunk = p.nchunk, pitch = p.pitch;
    const int stage_bytes = 16 * pitch;
    uint64_t *bars = (uint64_t *)smem;  // [S]
    float *red = (float *)(smem + 128); // [KS][16][Mp]
    uint8_t *wbuf = smem + 128 + (((size_t)KS * 16 * Mp * 4 + 127) & ~(size_t)127);

    const int T = p.N / 16;
    const int n_tiles = (int)blockIdx.x < T ? (T - 1 - (int)blockIdx.x) / (int)gridDim.x + 1 : 0;
    const int n_work = n_tiles * nchunk; // stages this CTA streams: (tile i, chunk c) = work / nchunk, work % nchunk
    auto issue = [&](int w, int slot) {
        const int i = w / nchunk, c = w % nchunk;
        const int row0 = ((int)blockIdx.x + i * (int)gridDim.x) * 16;
        const int sb0 = c * KC, nsbc = min(KC, nsb - sb0);
        const uint32_t bar = smem_u32(&bars[slot]);
        uint8_t *dst = wbuf + (size_t)slot * stage_bytes;
        mbar_expect_tx(bar, (uint32_t)(16 * nsbc * KQ_Q4K));
        for (int r = 0; r < 16; r++)
            bulk_g2s(smem_u32(dst + (size_t)r * pitch), p.W + ((size_t)(row0 + r) * nsb + sb0) * KQ_Q4K, (uint32_t)(nsbc * KQ_Q4K), bar);
    };
    if (tid == 0) {
        for (int s = 0; s < S; s++) mbar_init(smem_u32(&bars[s]), 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        for (int w = 0; w < S && w < n_work; w++) issue(w, w); // weights depend on nothing on the stream: ahead of the wait
    }
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    asm volatile("griddepcontrol.wait;" ::: "memory");
    __syncthreads();

    const int colw = mg * MTW * 8;
    // per batch tile (8 activation rows = one row group of the image): its cell row, clamped into the groups that exist (a tile past the
    // last row reads a valid group again; its columns are never stored)
    uint32_t cellb[MTW];
#pragma unroll
    for (int m = 0; m < MTW; m++) cellb[m] = (uint32_t)min(p.grp0 + (colw >> 3) + m, p.G - 1) * (uint32_t)nsb;
    const uint8_t *xq = p.xq;
    for (int i = 0; i < n_tiles; i++) {
        float acc[MTW][4]; // [0],[1]: weight row g, columns 2t, 2t+1; [2],[3]: row g + 8
#pragma unroll
        for (int m = 0; m < MTW; m++) acc[m][0] = acc[m][1] = acc[m][2] = acc[m][3] = 0.f;
        for (int c = 0; c < nchunk; c++) {
            const int w = i * nchunk + c, slot = w % S;
            mbar_wait(smem_u32(&bars[slot]), (uint32_t)((w / S) & 1));
            const uint8_t *st = wbuf + (size_t)slot * stage_bytes;
            const int sb0 = c * KC, nsbc = min(KC, nsb - sb0);
            for (int sl = ks; sl < nsbc; sl += KS) {
                const int sb = sb0 + sl;
                const uint8_t *r0 = st + (size_t)g * pitch + (size_t)sl * KQ_Q4K, *r1 = r0 + (size_t)8 * pitch;
                const uint32_t h0 = *(const uint32_t *)r0, h1 = *(const uint32_t *)r1; // d | dmin (fp16 pair)
                const float dw0 = __half2float(__ushort_as_half((unsigned short)(h0 & 0xFFFFu))), dm0 = __half2float(__ushort_as_half((unsigned short)(h0 >> 16)));
                const float dw1 = __half2float(__ushort_as_half((unsigned short)(h1 & 0xFFFFu))), dm1 = __half2float(__ushort_as_half((unsigned short)(h1 >> 16)));
                uint32_t sc0[2], mn0[2], sc1[2], mn1[2];
                q4k_scales8(*(const uint32_t *)(r0 + 4), *(const uint32_t *)(r0 + 8), *(const uint32_t *)(r0 + 12), sc0, mn0);
                q4k_scales8(*(const uint32_t *)(r1 + 4), *(const uint32_t *)(r1 + 8), *(const uint32_t *)(r1 + 12), sc1, mn1);
                int sumi[MTW][4];
#pragma unroll
                for (int m = 0; m < MTW; m++) sumi[m][0] = sumi[m][1] = sumi[m][2] = sumi[m][3] = 0;
#pragma unroll
                for (int pr = 0; pr < 4; pr++) { // sub-blocks 2 pr (low nibbles) and 2 pr + 1 (high nibbles) share their 32 bytes
                    // The contraction index of the MMA is only a label: lane t feeds k = 4t..4t+3 and 16+4t..16+4t+3 with elements 8t..8t+3 and
                    // 8t+4..8t+7 of the sub-block on BOTH operands, so every fragment is one 8-byte load (the activation gathers -- 8 rows, 8
                    // cache lines per instruction -- were the L1 pipe's main load: half as many now)
                    const uint2 wa = *(const uint2 *)(r0 + 16 + pr * 32 + 8 * t), wb = *(const uint2 *)(r1 + 16 + pr * 32 + 8 * t);
                    const uint32_t w00 = wa.x, w01 = wa.y, w10 = wb.x, w11 = wb.y;
#pragma unroll
                    for (int hi = 0; hi < 2; hi++) {
                        const int j = 2 * pr + hi;
                        const uint32_t a0 = (w00 >> (4 * hi)) & 0x0F0F0F0Fu, a1 = (w10 >> (4 * hi)) & 0x0F0F0F0Fu;
                        const uint32_t a2 = (w01 >> (4 * hi)) & 0x0F0F0F0Fu, a3 = (w11 >> (4 * hi)) & 0x0F0F0F0Fu;
                        const int s0 = byte_of(sc0, j), s1 = byte_of(sc1, j);
#pragma unroll
                        for (int m = 0; m < MTW; m++) {
                            const uint2 bb = *(const uint2 *)(xq + (size_t)(cellb[m] + sb) * 2048 + j * 256 + lane * 8); // 256 contiguous bytes per warp
                            int cc[4];
                            mma_u8s8(cc, a0, a1, a2, a3, bb.x, bb.y);
                            sumi[m][0] += s0 * cc[0]; sumi[m][1] += s0 * cc[1];
                            sumi[m][2] += s1 * cc[2]; sumi[m][3] += s1 * cc[3];
                        }
                    }
                }
                // The min term summ[row][col] = sum_j min_j[row] * bs_j[col], bs_j = bsums[2j] + bsums[2j+1] (|bs_j| <= 4096: 13 bits), as two more
                // MMAs instead of 8 multiply-adds per output: bs_j = 64 * hi_j + lo_j (hi_j = bs_j >> 6 in [-64, 63], lo_j in [0, 63]) are the B
                // operands (k = j, the other 24 k are zero), the eight 6-bit mins of the 16 rows the A operand; summ = 64 * S_hi + S_lo, exact.
                const uint32_t ma0 = t == 0 ? mn0[0] : t == 1 ? mn0[1] : 0u, ma1 = t == 0 ? mn1[0] : t == 1 ? mn1[1] : 0u; // k = 4t .. 4t+3 -> j
#pragma unroll
                for (int m = 0; m < MTW; m++) {
                    uint32_t bh = 0, bl = 0;
                    if (t < 2) { // B-fragment row = column (colw + 8 m + g): the sums of its sub-blocks j = 4t .. 4t+3
                        const uint2 gs = *(const uint2 *)(p.xbs + (size_t)(cellb[m] + sb) * 128 + g * 16 + t * 8);
                        const int bsv[4] = {(int)(short)(gs.x & 0xFFFFu), (int)(short)(gs.x >> 16), (int)(short)(gs.y & 0xFFFFu), (int)(short)(gs.y >> 16)};
#pragma unroll
                        for (int q = 0; q < 4; q++) {
                            bh |= ((uint32_t)(bsv[q] >> 6) & 0xFFu) << (8 * q);
                            bl |= ((uint32_t)bsv[q] & 63u) << (8 * q);
                        }
                    }
                    int sh[4], sl[4];
                    mma_u8s8(sh, ma0, ma1, 0u, 0u, bh, 0u);
                    mma_u8s8(sl, ma0, ma1, 0u, 0u, bl, 0u);
                    const float2 dxp = *(const float2 *)(p.xdx + (size_t)(cellb[m] + sb) * 32 + 8 * t); // d of columns 2t, 2t+1
                    const float dx0 = dxp.x, dx1 = dxp.y;
                    // the super-block's two fp32 FMAs (same operations as k_gemm_kq: d = d_x * d_w, -d_x * dmin_w)
                    acc[m][0] = fmaf(__fmul_rn(dx0, dw0), (float)sumi[m][0], acc[m][0]);
                    acc[m][1] = fmaf(__fmul_rn(dx1, dw0), (float)sumi[m][1], acc[m][1]);
                    acc[m][2] = fmaf(__fmul_rn(dx0, dw1), (float)sumi[m][2], acc[m][2]);
                    acc[m][3] = fmaf(__fmul_rn(dx1, dw1), (float)sumi[m][3], acc[m][3]);
                    acc[m][0] = fmaf(__fmul_rn(-dx0, dm0), (float)(64 * sh[0] + sl[0]), acc[m][0]);
                    acc[m][1] = fmaf(__fmul_rn(-dx1, dm0), (float)(64 * sh[1] + sl[1]), acc[m][1]);
                    acc[m][2] = fmaf(__fmul_rn(-dx0, dm1), (float)(64 * sh[2] + sl[2]), acc[m][2]);
                    acc[m][3] = fmaf(__fmul_rn(-dx1, dm1), (float)(64 * sh[3] + sl[3]), acc[m][3]);
                }
            }
            __syncthreads(); // every warp is done reading the slot
            if (tid == 0 && w + S < n_work) issue(w + S, slot);
        }
        // partial sums of the K-splits -> shared memory, added in a fixed order
        {
            float *r = red + ((size_t)ks * 16) * Mp;
#pragma unroll
            for (int m = 0; m < MTW; m++) {
                const int col0 = colw + m * 8 + 2 * t;
                *(float2 *)(r + (size_t)g * Mp + col0) = make_float2(acc[m][0], acc[m][1]);
                *(float2 *)(r + (size_t)(g + 8) * Mp + col0) = make_float2(acc[m][2], acc[m][3]);
            }
        }
        __syncthreads();
        const int row0 = ((int)blockIdx.x + i * (int)gridDim.x) * 16;
        for (int idx = tid; idx < 16 * p.M; idx += KQ_WARPS * 32) {
            const int r = idx & 15, m = idx >> 4;
            float v = 0.f;
#pragma unroll
            for (int k = 0; k < KS; k++) v += red[((size_t)k * 16 + r) * Mp + m];
            const int n = row0 + r;
            if (p.bias) v = __fadd_rn(v, p.bias[n]);
            const size_t at = (size_t)m * p.ldy + n;
            if (p.out_f16) ((__half *)p.y)[at] = __float2half_rn(v);
            else ((float *)p.y)[at] = v;
        }
        __syncthreads(); // red is reused by the next tile
    }
}

template <int MTW, int MG, int KS>
static int launch_q4k(const GemmQ4KParams &p, int grid, size_t smem, cudaStream_t st) {
    static size_t configured = 0;
    if (smem > configured) {
        CU_TRY(cudaFuncSetAttribute(k_gemm_mma_q4k<MTW, MG, KS>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        configured = smem;
    }
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(grid);
    cfg.blockDim = dim3(KQ_WARPS * 32);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    CU_TRY(cudaLaunchKernelEx(&cfg, k_gemm_mma_q4k<MTW, MG, KS>, p));
    return 0;
}

bool gemm_mma_q4k_supported(int M, int N, int K, const void *W) {
    static int on = -1, min_rows = 9;
    if (on < 0) {
        const char *e = getenv("MLLMCU_Q4K_MMA");
        on = (e && e[0] == '0') ? 0 : 1;
        const char *r = getenv("MLLMCU_Q4K_MMA_MIN");
        if (r && r[0]) min_rows = atoi(r);
    }
    return on == 1 && M >= min_rows && N > 0 && N % 16 == 0 && K > 0 && K % 256 == 0 && !((uintptr_t)W & 15);
}

// xq: the M activation rows as the fragment-major image (launch_quantize_q8_K_frag, kquants.cu).  Serves M in chunks of 64 rows (128 when K has fewer
// than 8 super-blocks): each chunk streams the weights again -- they stay in L2 between chunks.
int launch_gemm_mma_q4k(const void *W, const void *xq, const float *bias, void *y, int M, int N, int K, int ldy, int out_f16, cudaStream_t st) {
    if (!gemm_mma_q4k_supported(M, N, K, W)) return -2;
    const int nsb = K / 256;
    // row groups x K-splits of the 16 warps: up to 64 rows 2 x 8, up to 128 rows (or short K) 4 x 4; prefill-sized M gives every warp its own
    // 8 MTW rows over the whole K (16 x 1: no partial sums to merge, long steady loops), MTW = 2..5 -> 256..640 rows per pass over the weights
    static int big = -1;
    if (big < 0) { const char *e = getenv("MLLMCU_Q4K_MMA_BIG"); big = (e && e[0] == '0') ? 0 : 1; }
    int variant; // 0: <4,2,8>  1: <4,4,4>  2..5: <MTW,16,1>
    if (M > 128 && big) variant = M > 512 ? 5 : M > 384 ? 4 : M > 256 ? 3 : 2;
    else variant = (nsb < 8 || M > 64) ? 1 : 0;
    const int Mp = variant == 0 ? 64 : variant == 1 ? 128 : 128 * variant, KS = variant == 0 ? 8 : variant == 1 ? 4 : 1;
    GemmQ4KParams p = {};
    p.W = (const uint8_t *)W; p.bias = bias; p.N = N; p.nsb = nsb; p.ldy = ldy; p.out_f16 = out_f16;
    // a stage of 16 rows x KC super-blocks <= ~40 KB, the ring gets what the reduction buffer leaves
    int KC = nsb < 16 ? nsb : 16;
    p.KC = KC; p.nchunk = (nsb + KC - 1) / KC;
    p.pitch = KC * KQ_Q4K + 32; // 2304 + 32: rows g = 0..7 of a fragment start 8 banks apart (8-byte fragment loads: 4 rows x 32 bytes per half-warp)
    const size_t red = ((size_t)KS * 16 * Mp * 4 + 127) & ~(size_t)127, stage = (size_t)16 * p.pitch;
    int S = (int)((200 * 1024 - red) / stage);
    if (S > 8) S = 8;
    if (S < 2) return -2;
    p.stages = S;
    const size_t smem = 128 + red + (size_t)S * stage;
    int grid = cu_sm_count();
    if (grid > N / 16) grid = N / 16;
    const size_t esz = out_f16 ? 2 : 4;
    p.G = (M + 7) / 8;
    p.xq = (const uint8_t *)xq;
    p.xdx = p.xq + (size_t)p.G * nsb * 2048;
    p.xbs = p.xdx + (size_t)p.G * nsb * 32;
    for (int m0 = 0; m0 < M; m0 += Mp) {
        p.M = M - m0 < Mp ? M - m0 : Mp;
        p.grp0 = m0 / 8;
        p.y = (char *)y + (size_t)m0 * ldy * esz;
        int rc;
        switch (variant) {
        case 0: rc = launch_q4k<4, 2, 8>(p, grid, smem, st); break;
        case 1: rc = launch_q4k<4, 4, 4>(p, grid, smem, st); break;
        case 2: rc = launch_q4k<2, 16, 1>(p, grid, smem, st); break;
        case 3: rc = launch_q4k<3, 16, 1>(p, grid, smem, st); break;
        case 4: rc = launch_q4k<4, 16, 1>(p, grid, smem, st); break;
        default: rc = launch_q4k<5, 16, 1>(p, grid, smem, st); break;
        }
        if (rc) return rc;
    }
    return 0;
}
