// gemv_mma.cu -- batched decode Linear (3..64 activation rows) on the tensor cores.
//
// Reference arithmetic: the same as gemv_tma.cu -- mat_mul's generic path (compute/Matmul.cpp:77-120,219-276) with
// quantize_row_q8_0 (ggml/QuantizeQ8.cpp:106-160) and vec_dot_q4_0_q8_0 (ggml/VecDotQ4.cpp:514-544): per 32-element block the exact
// integer dot of the Q4_0 nibbles with the Q8_0 activations, times fp16(d_w) * fp16(d_x), accumulated in fp32.  This is the
// demo_qwen_batch shape (examples/demo_qwen_batch.cpp:61-68, Module::generate over a batch, Module.cpp:146-205): every decode step
// multiplies B rows with every weight matrix.
//
// Why a second kernel: k_gemv_tma does 8 dp4a per block PER BATCH ROW on the CUDA cores (21 warp instructions per 18-byte block at
// 8 rows), so the step time grew linearly with the batch (0.08 of the HBM roofline at batch 8).  With 8+ rows the product is a small
// dense contraction: here 16 weight rows x one Q4_0 block are the A operand (16 x 32, u8 nibbles as stored) and 8 batch rows x one Q8_0
// block the B operand (32 x 8, s8) of ONE mma.sync.m16n8k32.s8.s8 (nibbles made signed, nibble - 8, once per block and reused for
// every batch tile) -- its 128 int32 results are the exact block sums Sum((nibble - 8) * q8) of the reference.
//
// Layout: weights stay in the .mllm native layout; a tile of 16 consecutive rows is one contiguous byte range, staged by ONE
// cp.async.bulk (TMA) per matrix into a ring of shared-memory slots (filled before the programmatic-dependent-launch wait).  The
// activation image is made once per launch by k_act_quant_mma (RMSNorm + Q8_0, bytes identical to the stand-alone quantiser):
// q[block][row][32] int8 and ds[block][row] = (d, Sum q) -- block-major so the B fragment of 8 rows is one coalesced 256-byte read
// from L2.  16 warps per CTA = KS K-splits x MT batch tiles; a warp walks its blocks in order, partial sums of the K-splits meet in
// shared memory and are added in a fixed order (deterministic); epilogues as the GEMV: bias, + residual, silu(gate) * up.
#include "common.cuh"
#include "gemv_common.cuh"
#include "mllmcu.h"
#include <stdlib.h>

#define MMA_WARPS 16
#define MMA_MAX_STAGES 8

struct GemvMmaParams {
    const uint8_t *W[3];
    const float *bias[3];
    int N[3];
    int nW, K, M, Mp; // Mp: M rounded up to 8 (image rows)
    const int8_t *q;  // [nb][Mp][32]
    const float2 *ds; // [nb][Mp]
    const float *residual;
    int ldr;
    void *y;
    int ldy, out_f16, stages, late_trigger;
    int pitch; // bytes between consecutive weight rows of a tile in shared memory (row bytes, + 16 when those are a multiple of 128)
};

static __device__ __forceinline__ void mma_s8s8(int (&c)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t b0, uint32_t b1) {
    asm volatile("mma.sync.aligned.m16n8k32.row.col.s32.s8.s8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%10,%11,%12,%13};"
                 : "=r"(c[0]), "=r"(c[1]), "=r"(c[2]), "=r"(c[3])
                 : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1), "r"(0), "r"(0), "r"(0), "r"(0));
}

// four packed nibbles 0..15 -> four int8 (nibble - 8), the Q4_0 offset: +0x78 never carries out of a byte, the XOR flips bit 7
static __device__ __forceinline__ uint32_t nib_to_s8(uint32_t v) { return (v + 0x78787878u) ^ 0x80808080u; }

// 4 nibble bytes of block b of a weight row in shared memory, starting at element 4t (the row base is 16-byte aligned, the block
// 2-byte aligned): aligned word(s) + funnel shift
static __device__ __forceinline__ uint32_t ld_qs4(const uint8_t *row, int b, int t) {
    const uint32_t off = (uint32_t)b * 18u + 2u + 4u * (uint32_t)t;
    const uint32_t *wp = (const uint32_t *)(row + (off & ~3u));
    if (off & 2u) return __byte_perm(wp[0], wp[1], 0x5432);
    return wp[0];
}

// MTW batch tiles (8 rows each) per warp, MG groups of them, KS K-splits: MG * KS = 16 warps, Mp = MG * MTW * 8.
// A warp unpacks the A fragment of a block ONCE and reuses it for its MTW batch tiles; the B fragments and scales of all MTW tiles
// are requested before the first MMA so their L2 latency overlaps (2 blocks in flight through the unroll).
template <int MTW, int MG, int KS, int EPI>
__global__ void __launch_bounds__(MMA_WARPS * 32) k_gemv_mma(const GemvMmaParams p) {
    static_assert(MG * KS == MMA_WARPS, "16 warps");
    extern __shared__ __align__(128) uint8_t smem[];
    constexpr int NWT = (EPI == 2) ? 2 : 1;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int g = lane >> 2, t = lane & 3;
    const int ks = warp % KS, mg = warp / KS;
    const int K = p.K, nb = K / QK, rowb = nb * 18, S = p.stages, Mp = p.Mp;
    // Rows whose byte length is a multiple of 128 (K = 2048, 4096 ...) would put rows g = 0..7 of the A fragment in the same banks
    // (8-way conflicts on every LDS): such tiles are staged row by row with a 16-byte pad, so row g starts 16 g bytes into the banks
    const int pitch = p.pitch;
    const int tile_bytes = 16 * pitch, stage_bytes = tile_bytes * NWT;
    uint64_t *bars = (uint64_t *)smem;                 // [S]
    float *red = (float *)(smem + 128);                // [NWT][KS][16][Mp]
    uint8_t *wbuf = smem + 128 + (((size_t)NWT * KS * 16 * Mp * 4 + 127) & ~(size_t)127);

    int T = p.N[0] / 16;
    if (EPI != 2) for (int i = 1; i < p.nW; i++) T += p.N[i] / 16;
    const int n_my = (int)blockIdx.x < T ? (T - 1 - (int)blockIdx.x) / (int)gridDim.x + 1 : 0;
    auto locate = [&](int tile, int &mi, int &row0) { // tile -> (matrix, first row inside it)
        mi = 0;
        int tl = tile;
        if (EPI != 2) while (mi < p.nW - 1 && tl >= p.N[mi] / 16) { tl -= p.N[mi] / 16; mi++; }
        row0 = tl * 16;
    };
    auto issue = [&](int i, int slot) {
        int mi, row0;
        locate((int)blockIdx.x + i * (int)gridDim.x, mi, row0);
        const uint32_t bar = smem_u32(&bars[slot]);
        uint8_t *dst = wbuf + (size_t)slot * stage_bytes;
        mbar_expect_tx(bar, (uint32_t)(16 * rowb * NWT));
        if (pitch == rowb) {
            bulk_g2s(smem_u32(dst), p.W[mi] + (size_t)row0 * rowb, (uint32_t)(16 * rowb), bar);
            if (EPI == 2) bulk_g2s(smem_u32(dst + tile_bytes), p.W[1] + (size_t)row0 * rowb, (uint32_t)(16 * rowb), bar);
        } else {
            for (int r = 0; r < 16; r++) {
                bulk_g2s(smem_u32(dst + (size_t)r * pitch), p.W[mi] + (size_t)(row0 + r) * rowb, (uint32_t)rowb, bar);
                if (EPI == 2) bulk_g2s(smem_u32(dst + tile_bytes + (size_t)r * pitch), p.W[1] + (size_t)(row0 + r) * rowb, (uint32_t)rowb, bar);
            }
        }
    };
    if (tid == 0) {
        for (int s = 0; s < S; s++) mbar_init(smem_u32(&bars[s]), 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        for (int i = 0; i < S && i < n_my; i++) issue(i, i); // weights depend on nothing on the stream: ahead of the wait
    }
    if (!p.late_trigger) asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    asm volatile("griddepcontrol.wait;" ::: "memory");
    if (p.late_trigger) asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    __syncthreads(); // barrier inits visible to every waiter

    const int colw = mg * MTW * 8; // first batch row of this warp's tiles
    for (int i = 0; i < n_my; i++) {
        const int slot = i % S;
        mbar_wait(smem_u32(&bars[slot]), (uint32_t)((i / S) & 1));
        const uint8_t *w0 = wbuf + (size_t)slot * stage_bytes + (size_t)g * pitch; // row g of the tile; row g + 8 is 8 * pitch further
        float2 acc[NWT][MTW][2]; // [0]: row g, batch rows (2t, 2t+1) of the tile; [1]: row g + 8
#pragma unroll
        for (int j = 0; j < NWT; j++)
#pragma unroll
            for (int m = 0; m < MTW; m++) acc[j][m][0] = acc[j][m][1] = make_float2(0.f, 0.f);
#pragma unroll 2
        for (int b = ks; b < nb; b += KS) {
            // B fragments and scales of block b for this warp's batch tiles: coalesced L2 reads, all requested up front
            uint32_t b0[MTW], b1[MTW];
            float4 dsx[MTW];
#pragma unroll
            for (int m = 0; m < MTW; m++) {
                const int8_t *qb = p.q + ((size_t)b * Mp + colw + m * 8 + g) * 32;
                b0[m] = *(const uint32_t *)(qb + 4 * t);
                b1[m] = *(const uint32_t *)(qb + 16 + 4 * t);
                dsx[m] = *(const float4 *)(p.ds + (size_t)b * Mp + colw + m * 8 + 2 * t); // (d, sum) of the thread's two batch rows
            }
#pragma unroll
            for (int j = 0; j < NWT; j++) {
                const uint8_t *r0 = w0 + (size_t)j * tile_bytes, *r1 = r0 + (size_t)8 * pitch;
                const uint32_t q0 = ld_qs4(r0, b, t), q1 = ld_qs4(r1, b, t);
                const float dw0 = __half2float(*(const __half *)(r0 + (size_t)b * 18)), dw1 = __half2float(*(const __half *)(r1 + (size_t)b * 18));
                // signed once per block (amortised over the MTW batch tiles), so the MMA returns Sum((nibble - 8) * q8) itself
                const uint32_t a0 = nib_to_s8(q0 & 0x0F0F0F0Fu), a1 = nib_to_s8(q1 & 0x0F0F0F0Fu);
                const uint32_t a2 = nib_to_s8((q0 >> 4) & 0x0F0F0F0Fu), a3 = nib_to_s8((q1 >> 4) & 0x0F0F0F0Fu);
                const float2 dw0_2 = make_float2(dw0, dw0), dw1_2 = make_float2(dw1, dw1);
#pragma unroll
                for (int m = 0; m < MTW; m++) {
                    int c[4];
                    mma_s8s8(c, a0, a1, a2, a3, b0[m], b1[m]);
                    // packed fp32x2: (d_w * d_x) rounded as the reference's product, then one FMA per element (VecDotQ4.cpp:523-543)
                    const float2 dx2 = make_float2(dsx[m].x, dsx[m].z);
                    acc[j][m][0] = __ffma2_rn(__fmul2_rn(dw0_2, dx2), make_float2((float)c[0], (float)c[1]), acc[j][m][0]);
                    acc[j][m][1] = __ffma2_rn(__fmul2_rn(dw1_2, dx2), make_float2((float)c[2], (float)c[3]), acc[j][m][1]);
                }
            }
        }
        // partial sums of the K-splits -> shared memory
#pragma unroll
        for (int j = 0; j < NWT; j++) {
            float *r = red + ((size_t)(j * KS + ks) * 16) * Mp;
#pragma unroll
            for (int m = 0; m < MTW; m++) {
                const int col0 = colw + m * 8 + 2 * t;
                *(float2 *)(r + (size_t)g * Mp + col0) = acc[j][m][0];
                *(float2 *)(r + (size_t)(g + 8) * Mp + col0) = acc[j][m][1];
            }
        }
        __syncthreads(); // every warp is also done reading the slot
        if (tid == 0 && i + S < n_my) issue(i + S, slot);
        int mi, row0;
        locate((int)blockIdx.x + i * (int)gridDim.x, mi, row0);
        int colbase = row0; // column of the tile's first row in the (concatenated) output
        if (EPI != 2) for (int k = 0; k < mi; k++) colbase += p.N[k];
        for (int idx = tid; idx < 16 * p.M; idx += MMA_WARPS * 32) {
            const int r = idx & 15, m = idx >> 4;
            float v = 0.f, v2 = 0.f;
#pragma unroll
            for (int k = 0; k < KS; k++) v += red[((size_t)k * 16 + r) * Mp + m]; // fixed order: deterministic
            if (EPI == 2) {
#pragma unroll
                for (int k = 0; k < KS; k++) v2 += red[((size_t)(KS + k) * 16 + r) * Mp + m];
            }
            const int nloc = row0 + r, ncol = colbase + r;
            if (p.bias[mi]) v = v + p.bias[mi][nloc];
            if (EPI == 1) v = __fadd_rn(v, p.residual[(size_t)m * p.ldr + ncol]);
            if (EPI == 2) {
                if (p.bias[1]) v2 = v2 + p.bias[1][nloc];
                v = __fmul_rn(g_silu(v), v2);
            }
            const size_t at = (size_t)m * p.ldy + ncol;
            if (p.out_f16) ((__half *)p.y)[at] = __float2half_rn(v);
            else ((float *)p.y)[at] = v;
        }
        __syncthreads(); // red is reused by the next tile
    }
}

// ------------------------------------------------------------------ activation image
// (RMSNorm +) Q8_0 quantisation of the M rows, one CTA per row, written block-major for the MMA kernel.  Same arithmetic as
// k_act_quant / quantize.cu (bytes identical; checked through mllmcu_gemv_fused against the oracle).
__global__ void __launch_bounds__(256) k_act_quant_mma(const float *__restrict__ x, int ldx, const float *__restrict__ norm_w, float eps, int K, int M, int Mp,
                                                       int8_t *__restrict__ q, float2 *__restrict__ ds) {
    __shared__ double red[8];
    __shared__ float s_rms;
    const int m = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int nb = K / QK;
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); // the GEMV may start and fill its weight ring
    asm volatile("griddepcontrol.wait;" ::: "memory");
    if (m >= M) { // padding rows of the last batch tile: zeros
        for (int b = tid; b < nb; b += blockDim.x) {
            uint4 *d = (uint4 *)(q + ((size_t)b * Mp + m) * 32);
            d[0] = make_uint4(0, 0, 0, 0);
            d[1] = make_uint4(0, 0, 0, 0);
            ds[(size_t)b * Mp + m] = make_float2(0.f, 0.f);
        }
        return;
    }
    const float *xr = x + (size_t)m * ldx;
    float rms = 1.0f;
    if (norm_w) { // CPURMSNorm.cpp:45-62: sum of squares in double, mean in float
        double a = 0.0;
        for (int d = tid * 4; d < K; d += blockDim.x * 4) {
            const float4 v = *(const float4 *)(xr + d);
            a = fma((double)v.x, (double)v.x, a); a = fma((double)v.y, (double)v.y, a);
            a = fma((double)v.z, (double)v.z, a); a = fma((double)v.w, (double)v.w, a);
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) a += __shfl_xor_sync(0xffffffffu, a, o);
        if (lane == 0) red[warp] = a;
        __syncthreads();
        if (tid == 0) {
            double tot = 0.0;
            for (int w = 0; w < (int)(blockDim.x >> 5); w++) tot += red[w];
            const float mean = (float)(tot / (double)K);
            s_rms = __fdiv_rn(1.0f, __fsqrt_rn(__fadd_rn(mean, eps)));
        }
        __syncthreads();
        rms = s_rms;
    }
    // 8 lanes per block: lane `sub` holds elements [4 sub, 4 sub + 4)
    const int sub = lane & 7, grp = lane >> 3;
    for (int b0 = warp * 4; b0 < nb; b0 += (blockDim.x >> 5) * 4) { // warp-uniform trip count (shuffles inside)
        const int b = b0 + grp;
        const bool ok = b < nb;
        float4 v = ok ? *(const float4 *)(xr + (size_t)b * QK + sub * 4) : make_float4(0.f, 0.f, 0.f, 0.f);
        if (norm_w && ok) {
            const float4 w = *(const float4 *)(norm_w + (size_t)b * QK + sub * 4);
            v.x = __fmul_rn(__fmul_rn(v.x, rms), w.x); v.y = __fmul_rn(__fmul_rn(v.y, rms), w.y);
            v.z = __fmul_rn(__fmul_rn(v.z, rms), w.z); v.w = __fmul_rn(__fmul_rn(v.w, rms), w.w);
        }
        float amax = fmaxf(fmaxf(fabsf(v.x), fabsf(v.y)), fmaxf(fabsf(v.z), fabsf(v.w)));
        amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, 1));
        amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, 2));
        amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, 4));
        float id, d;
        q8_0_id_d(amax, id, d);
        const int q0 = __float2int_rn(__fmul_rn(v.x, id)), q1 = __float2int_rn(__fmul_rn(v.y, id));
        const int q2 = __float2int_rn(__fmul_rn(v.z, id)), q3 = __float2int_rn(__fmul_rn(v.w, id));
        int sum = q0 + q1 + q2 + q3;
        sum += __shfl_xor_sync(0xffffffffu, sum, 1);
        sum += __shfl_xor_sync(0xffffffffu, sum, 2);
        sum += __shfl_xor_sync(0xffffffffu, sum, 4);
        if (ok) {
            const uint32_t packed = __byte_perm(__byte_perm((uint32_t)q0, (uint32_t)q1, 0x0040), __byte_perm((uint32_t)q2, (uint32_t)q3, 0x0040), 0x5410);
            *(uint32_t *)(q + ((size_t)b * Mp + m) * 32 + sub * 4) = packed;
            if (sub == 0) ds[(size_t)b * Mp + m] = make_float2(__half2float(__float2half_rn(d)), __int_as_float(sum));
        }
    }
}

// ------------------------------------------------------------------ host side
template <int MTW, int MG, int KS, int EPI>
static int launch_mma(const GemvMmaParams &p, int grid, size_t smem, cudaStream_t st, int pdl) {
    static size_t configured = 0;
    if (smem > configured) {
        CU_TRY(cudaFuncSetAttribute(k_gemv_mma<MTW, MG, KS, EPI>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        configured = smem;
    }
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(grid);
    cfg.blockDim = dim3(MMA_WARPS * 32);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = pdl ? 1 : 0;
    CU_TRY(cudaLaunchKernelEx(&cfg, k_gemv_mma<MTW, MG, KS, EPI>, p));
    return 0;
}
// batch tiles -> (tiles per warp, groups, K-splits)
static int mma_ksplits(int mtiles) { return mtiles == 8 ? 8 : 16; }
template <int EPI>
static int launch_mma_m(int mtiles, const GemvMmaParams &p, int grid, size_t smem, cudaStream_t st, int pdl) {
    switch (mtiles) {
    case 1: return launch_mma<1, 1, 16, EPI>(p, grid, smem, st, pdl);
    case 2: return launch_mma<2, 1, 16, EPI>(p, grid, smem, st, pdl);
    case 4: return launch_mma<4, 1, 16, EPI>(p, grid, smem, st, pdl);
    case 8: return launch_mma<4, 2, 8, EPI>(p, grid, smem, st, pdl);
    default: return -2;
    }
}

// scratch for the activation image: per device, grown on demand outside stream capture (64 rows x K 16384 = 1.3 MB at most)
static void *g_img[64] = {};
static size_t g_img_bytes[64] = {};

// rows from which the tensor-core kernel wins over the CUDA-core GEMV (Qwen1.5-1.8B step: 4 rows 1.41 vs 1.33 ms, 8 rows 1.30 vs 1.98)
int gemv_mma_min_rows() {
    static int v = -1;
    if (v < 0) { const char *e = getenv("MLLMCU_GEMV_MMA_MIN"); v = (e && e[0]) ? atoi(e) : 5; if (v < 3) v = 3; }
    return v;
}
bool gemv_mma_supported(int wtype, int M, int K, const int *Ns, int nW, const void *const *W) {
    if (wtype != 2 || M < gemv_mma_min_rows() || M > 64 || K <= 0 || K % 128 || K > 16384) return false;
    for (int i = 0; i < nW; i++)
        if (Ns[i] <= 0 || Ns[i] % 16 || ((uintptr_t)W[i] & 15)) return false;
    return true;
}

// The batched decode Linear behind mllmcu_gemv_fused / _ex for 3 <= M <= 64 (same argument meaning).  Returns -2 when the shape is
// outside this kernel (caller falls back to k_gemv_tma for M <= 8, else to mllmcu_linear).
int launch_gemv_mma(const mllmcu_gemv_desc &D, void *stream) {
    const int M = D.M, K = D.K, nW = D.nW, epi = D.epilogue;
    if (!gemv_mma_supported(D.wtype, M, K, D.N, nW, D.W)) return -2;
    if (epi == 2 && (nW != 2 || D.N[0] != D.N[1])) return -1;
    if (epi == 1 && D.residual == nullptr) return -1;
    if (((uintptr_t)D.x & 15) || (D.ldx & 3) || (D.norm_w && ((uintptr_t)D.norm_w & 15))) return -2;
    cudaStream_t st = (cudaStream_t)stream;
    int dev = 0;
    CU_TRY(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64) return -1;
    static int sm_count[64] = {};
    if (sm_count[dev] == 0) CU_TRY(cudaDeviceGetAttribute(&sm_count[dev], cudaDevAttrMultiProcessorCount, dev));
    // batch tiles: 1, 2, 4 or 8 (rows padded up; padded rows are zeros in the image and never stored)
    int mtiles = (M + 7) / 8;
    if (mtiles == 3) mtiles = 4;
    if (mtiles > 4) mtiles = 8;
    const int Mp = mtiles * 8, nb = K / QK, ks = mma_ksplits(mtiles);
    const size_t img = (size_t)nb * Mp * 32 + (size_t)nb * Mp * 8 + 256;
    if (img > g_img_bytes[dev]) {
        cudaStreamCaptureStatus cs = cudaStreamCaptureStatusNone;
        CU_TRY(cudaStreamIsCapturing(st, &cs));
        if (cs != cudaStreamCaptureStatusNone) return -2; // never allocate under capture: the first (uncaptured) step sizes the scratch
        const size_t want = (size_t)(16384 / QK) * 64 * 40 + 256; // the largest image this kernel takes
        CU_TRY(cudaStreamSynchronize(st));
        if (g_img[dev]) CU_TRY(cudaFree(g_img[dev]));
        CU_TRY(cudaMalloc(&g_img[dev], want));
        g_img_bytes[dev] = want;
    }
    int8_t *q = (int8_t *)g_img[dev];
    float2 *ds = (float2 *)((char *)g_img[dev] + (((size_t)nb * Mp * 32 + 255) & ~(size_t)255));
    const int pdl = D.pdl;
    {
        cudaLaunchConfig_t qc = {};
        qc.gridDim = dim3(Mp);
        qc.blockDim = dim3(256);
        qc.stream = st;
        cudaLaunchAttribute qa[1];
        qa[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        qa[0].val.programmaticStreamSerializationAllowed = 1;
        qc.attrs = qa;
        qc.numAttrs = pdl ? 1 : 0;
        CU_TRY(cudaLaunchKernelEx(&qc, k_act_quant_mma, D.x, D.ldx, D.norm_w, D.eps, K, M, Mp, q, ds));
    }
    GemvMmaParams p = {};
    int T = 0;
    for (int i = 0; i < nW; i++) { p.W[i] = (const uint8_t *)D.W[i]; p.bias[i] = D.bias[i]; p.N[i] = D.N[i]; T += D.N[i] / 16; }
    if (epi == 2) T = D.N[0] / 16;
    p.nW = nW; p.K = K; p.M = M; p.Mp = Mp; p.q = q; p.ds = ds; p.residual = D.residual; p.ldr = D.ldr; p.y = D.y; p.ldy = D.ldy; p.out_f16 = D.out_f16;
    p.late_trigger = (pdl == 2) ? 1 : 0;
    const int nwt = epi == 2 ? 2 : 1;
    const int rowb = nb * 18;
    p.pitch = (rowb % 128 == 0) ? rowb + 16 : rowb;
    const size_t stage = (size_t)16 * p.pitch * nwt;
    const size_t fixed = 128 + (((size_t)nwt * ks * 16 * Mp * 4 + 127) & ~(size_t)127);
    // two CTAs per SM when the ring fits in half the shared memory with >= 2 slots, else one
    int stages = (int)((110 * 1024 - fixed) / stage);
    int per_sm = 2;
    if (stages < 2) { stages = (int)((225 * 1024 - fixed) / stage); per_sm = 1; }
    if (stages < 1) return -2;
    if (stages > MMA_MAX_STAGES) stages = MMA_MAX_STAGES;
    p.stages = stages;
    int grid = sm_count[dev] * per_sm;
    if (grid > T) grid = T;
    const size_t smem = fixed + (size_t)stages * stage;
    if (epi == 0) return launch_mma_m<0>(mtiles, p, grid, smem, st, pdl);
    if (epi == 1) return launch_mma_m<1>(mtiles, p, grid, smem, st, pdl);
    return launch_mma_m<2>(mtiles, p, grid, smem, st, pdl);
}
