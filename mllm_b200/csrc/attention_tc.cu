// attention_tc.cu -- prefill / append attention on the tensor cores.
// Reference: flash_attention_2_forward (compute/FlashAttention2.hpp:2236-2284), prefill-append body
// (:1212-1560 fp16 KV, :87-700 fp32 KV): fp32 Q, fp16 (or fp32) K/V, fp32 scores and accumulators,
// p = expf((s - rowmax) / sqrt(D)), o = acc / sum(p); key c visible to query r iff c <= r + (Sk - Sq).
//
// The reference multiplies fp32 q by the widened fp16 k in fp32.  To keep that precision on fp16 tensor
// cores every fp32 operand is split into two halves, x = hi + lo with hi = fp16(x), lo = fp16(x - hi):
//   s  = q_hi.k + q_lo.k          (k is fp16 already: both products are exact in the fp32 accumulator)
//   o += p_hi.v + p_lo.v          (p in [0,1])
// so the residual is |x| * 2^-22 per operand -- the size of an fp32 rounding, far inside the 1e-5 bar the parity
// tests hold.  fp32 K/V (vision towers, no cache) are split the same way at tile load: three products per
// contraction (hi.hi + lo.hi + hi.lo).
//
// fp32 K/V over long sequences (Qwen2-VL's tower: 5476 patches) are split ONCE per call by k_split_kv into four fp16 planes (K hi, K lo, V hi,
// V lo) in the caller's workspace; the attention kernel then streams those planes with cp.async, two stages, exactly like fp16 K/V (MODE 2)
// instead of every 64-row CTA re-reading and re-splitting all of K and V between two barriers (MODE 1, kept for callers without workspace).
//
// One CTA = 4 warps x 16 query rows of one (batch, head); keys in tiles of 64, staged in shared memory
// (fp16 K/V: cp.async, two stages; 16-byte chunks XOR-swizzled by row so ldmatrix is conflict-free), fragments through
// ldmatrix (K as stored, V transposed), mma.sync.m16n8k16 fp16 -> fp32, online softmax in registers (FA-2 layout).
// Work: 4 * Sq * Sk/2 * Hq * D flop per layer (causal), x2 for the split; the old warp-per-row CUDA-core kernel
// (attention.cu) remains for short query runs (Sq < 16) and head dims outside {64, 80, 128}.
#include "common.cuh"
#include <stdlib.h>

static __device__ __forceinline__ void mma16816(float (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
                 : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
static __device__ __forceinline__ void ldsm4(uint32_t (&r)[4], uint32_t addr) {
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];" : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(addr));
}
static __device__ __forceinline__ void ldsm4t(uint32_t (&r)[4], uint32_t addr) {
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];" : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(addr));
}
static __device__ __forceinline__ void cp_async16(uint32_t dst, const void *src) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst), "l"(src));
}
static __device__ __forceinline__ void split2(float x, float y, uint32_t &hi, uint32_t &lo) {
    const __half2 h = __floats2half2_rn(x, y);
    const float2 hf = __half22float2(h);
    const __half2 l = __floats2half2_rn(x - hf.x, y - hf.y);
    hi = *(const uint32_t *)&h;
    lo = *(const uint32_t *)&l;
}

// D: head dim (multiple of 16).  MODE 0: fp16 K/V.  MODE 1: fp32 K/V, split into hi/lo tiles at load.  MODE 2: fp32 K/V already split into
// fp16 planes (kp = K hi; K lo, V hi, V lo follow at multiples of `pstride` halves; vp unused).
// Shared-memory rows: MODE 0/1 a power-of-two number of 16-byte chunks >= D/8 with an XOR swizzle by row; MODE 2 D/8 + 1 chunks, no swizzle
// (row pitch = 4 banks mod 32: the 8 rows of an ldmatrix phase land in 8 different bank groups, and D = 80 needs 176 bytes instead of 256).
template <int D, int MODE>
__global__ void __launch_bounds__(128) k_attn_prefill_tc(const float *__restrict__ q, const void *__restrict__ kp, const void *__restrict__ vp,
                                                         float *__restrict__ o, int Sq, int Sk, int Hq, int Hkv, int causal, long long kvb,
                                                         long long pstride) {
    constexpr bool F32KV = MODE != 0, PAD = MODE == 2;
    constexpr int BC = 64;           // keys per tile
    constexpr int KS = D / 16;       // k-steps of q.k
    constexpr int NB = BC / 8;       // score column blocks
    constexpr int DB = D / 8;        // output column blocks
    constexpr int CH = D / 8;        // 16-byte chunks per row that hold data
    constexpr int RCH = PAD ? CH + 1 : (CH <= 8) ? 8 : 16; // chunks per shared-memory row
    constexpr int RB = RCH * 16;     // row bytes
    constexpr int TILE = BC * RB;    // bytes of one tile
    constexpr int NPL = F32KV ? 2 : 1;      // planes per operand (hi, lo)
    extern __shared__ __align__(128) unsigned char smem[];
    // [stage][K planes | V planes]
    const uint32_t sbase = (uint32_t)__cvta_generic_to_shared(smem);
    auto soff = [](int r, int c) { return (uint32_t)(PAD ? r * RB + (c << 4) : r * RB + ((c ^ (r & 7)) << 4)); };
    auto k_tile = [&](int st, int pl) { return sbase + (uint32_t)((st * 2 * NPL + pl) * TILE); };
    auto v_tile = [&](int st, int pl) { return sbase + (uint32_t)((st * 2 * NPL + NPL + pl) * TILE); };

    const int tid = threadIdx.x, w = tid >> 5, lane = tid & 31, g = lane >> 2, t = lane & 3;
    const int qt = gridDim.x - 1 - blockIdx.x; // longest (most keys) tiles first
    const int h = blockIdx.y, b = blockIdx.z;
    const int kvh = h / (Hq / Hkv);
    const int off = Sk - Sq;
    const int row0 = qt * 64 + w * 16;
    const float scale = __fdiv_rn(1.0f, __fsqrt_rn((float)D));
    const size_t rs = (size_t)Hkv * D;

    // ---- Q fragments (hi / lo), straight from global
    uint32_t qh[KS][4], ql[KS][4];
#pragma unroll
    for (int kk = 0; kk < KS; kk++) {
#pragma unroll
        for (int i = 0; i < 4; i++) {
            const int r = row0 + g + (i & 1) * 8;
            const int c = kk * 16 + 2 * t + (i >> 1) * 8;
            float2 x = make_float2(0.f, 0.f);
            if (r < Sq) x = *(const float2 *)(q + (((size_t)b * Sq + r) * Hq + h) * D + c);
            split2(x.x, x.y, qh[kk][i], ql[kk][i]);
        }
    }

    const int last = causal ? min(Sk - 1, qt * 64 + 63 + off) : (Sk - 1);
    const int ntile = last / BC + 1;

    auto load_tile = [&](int jt, int st) {
        if (MODE == 2) {
            const __half *pb = (const __half *)kp + (size_t)b * kvb + (size_t)kvh * D;
            for (int i = tid; i < BC * CH; i += 128) {
                const int r = i / CH, c = i % CH;
                const int key = jt * BC + r;
                const uint32_t so = soff(r, c);
                if (key < Sk) {
                    const __half *src = pb + (size_t)key * rs + c * 8;
                    cp_async16(k_tile(st, 0) + so, src);
                    cp_async16(k_tile(st, 1) + so, src + pstride);
                    cp_async16(v_tile(st, 0) + so, src + 2 * pstride);
                    cp_async16(v_tile(st, 1) + so, src + 3 * pstride);
                } else {
                    asm volatile("st.shared.v4.u32 [%0], {%1,%1,%1,%1};" ::"r"(k_tile(st, 0) + so), "r"(0));
                    asm volatile("st.shared.v4.u32 [%0], {%1,%1,%1,%1};" ::"r"(k_tile(st, 1) + so), "r"(0));
                    asm volatile("st.shared.v4.u32 [%0], {%1,%1,%1,%1};" ::"r"(v_tile(st, 0) + so), "r"(0));
                    asm volatile("st.shared.v4.u32 [%0], {%1,%1,%1,%1};" ::"r"(v_tile(st, 1) + so), "r"(0));
                }
            }
            asm volatile("cp.async.commit_group;");
        } else if (MODE == 0) {
            const __half *kb = (const __half *)kp + (size_t)b * kvb + (size_t)kvh * D;
            const __half *vb = (const __half *)vp + (size_t)b * kvb + (size_t)kvh * D;
            for (int i = tid; i < BC * CH; i += 128) {
                const int r = i / CH, c = i % CH;
                const int key = jt * BC + r;
                const uint32_t so = soff(r, c);
                if (key < Sk) {
                    cp_async16(k_tile(st, 0) + so, kb + (size_t)key * rs + c * 8);
                    cp_async16(v_tile(st, 0) + so, vb + (size_t)key * rs + c * 8);
                } else {
                    asm volatile("st.shared.v4.u32 [%0], {%1,%1,%1,%1};" ::"r"(k_tile(st, 0) + so), "r"(0));
                    asm volatile("st.shared.v4.u32 [%0], {%1,%1,%1,%1};" ::"r"(v_tile(st, 0) + so), "r"(0));
                }
            }
            asm volatile("cp.async.commit_group;");
        } else {
            const float *kb = (const float *)kp + (size_t)b * kvb + (size_t)kvh * D;
            const float *vb = (const float *)vp + (size_t)b * kvb + (size_t)kvh * D;
            for (int i = tid; i < BC * CH; i += 128) {
                const int r = i / CH, c = i % CH;
                const int key = jt * BC + r;
                const uint32_t so = soff(r, c);
                uint32_t kh4[4] = {0, 0, 0, 0}, kl4[4] = {0, 0, 0, 0}, vh4[4] = {0, 0, 0, 0}, vl4[4] = {0, 0, 0, 0};
                if (key < Sk) {
                    const float4 k0 = *(const float4 *)(kb + (size_t)key * rs + c * 8), k1 = *(const float4 *)(kb + (size_t)key * rs + c * 8 + 4);
                    const float4 v0 = *(const float4 *)(vb + (size_t)key * rs + c * 8), v1 = *(const float4 *)(vb + (size_t)key * rs + c * 8 + 4);
                    split2(k0.x, k0.y, kh4[0], kl4[0]); split2(k0.z, k0.w, kh4[1], kl4[1]);
                    split2(k1.x, k1.y, kh4[2], kl4[2]); split2(k1.z, k1.w, kh4[3], kl4[3]);
                    split2(v0.x, v0.y, vh4[0], vl4[0]); split2(v0.z, v0.w, vh4[1], vl4[1]);
                    split2(v1.x, v1.y, vh4[2], vl4[2]); split2(v1.z, v1.w, vh4[3], vl4[3]);
                }
                asm volatile("st.shared.v4.u32 [%0], {%1,%2,%3,%4};" ::"r"(k_tile(st, 0) + so), "r"(kh4[0]), "r"(kh4[1]), "r"(kh4[2]), "r"(kh4[3]));
                asm volatile("st.shared.v4.u32 [%0], {%1,%2,%3,%4};" ::"r"(k_tile(st, NPL - 1) + so), "r"(kl4[0]), "r"(kl4[1]), "r"(kl4[2]), "r"(kl4[3]));
                asm volatile("st.shared.v4.u32 [%0], {%1,%2,%3,%4};" ::"r"(v_tile(st, 0) + so), "r"(vh4[0]), "r"(vh4[1]), "r"(vh4[2]), "r"(vh4[3]));
                asm volatile("st.shared.v4.u32 [%0], {%1,%2,%3,%4};" ::"r"(v_tile(st, NPL - 1) + so), "r"(vl4[0]), "r"(vl4[1]), "r"(vl4[2]), "r"(vl4[3]));
            }
        }
    };

    float acc[DB][4];
#pragma unroll
    for (int d = 0; d < DB; d++) acc[d][0] = acc[d][1] = acc[d][2] = acc[d][3] = 0.f;
    float m0 = -INFINITY, m1 = -INFINITY, l0 = 0.f, l1 = 0.f; // rows g and g+8

    constexpr bool ASYNC = MODE != 1;
    if (ASYNC) load_tile(0, 0);
    for (int jt = 0; jt < ntile; jt++) {
        const int st = ASYNC ? (jt & 1) : 0;
        if (ASYNC) {
            asm volatile("cp.async.wait_group 0;");
            __syncthreads();
            if (jt + 1 < ntile) load_tile(jt + 1, st ^ 1);
        } else {
            __syncthreads(); // everyone is done with the previous tile
            load_tile(jt, 0);
            __syncthreads();
        }

        // ---- scores: S[16 x 64] = Q K^T
        float s[NB][4];
#pragma unroll
        for (int nb = 0; nb < NB; nb++) s[nb][0] = s[nb][1] = s[nb][2] = s[nb][3] = 0.f;
        const int mat = lane >> 3, mr = lane & 7;
#pragma unroll
        for (int nb = 0; nb < NB; nb++) {
            const int r = nb * 8 + mr;
#pragma unroll
            for (int kk2 = 0; kk2 < (KS + 1) / 2; kk2++) {
                const int c = 4 * kk2 + mat; // for odd KS the last pair's upper half reads pad chunks (unused)
                const uint32_t so = soff(r, c);
                uint32_t kf[4];
                ldsm4(kf, k_tile(st, 0) + so);
                mma16816(s[nb], qh[2 * kk2], kf[0], kf[1]);
                mma16816(s[nb], ql[2 * kk2], kf[0], kf[1]);
                if (2 * kk2 + 1 < KS) {
                    mma16816(s[nb], qh[2 * kk2 + 1], kf[2], kf[3]);
                    mma16816(s[nb], ql[2 * kk2 + 1], kf[2], kf[3]);
                }
                if (F32KV) {
                    uint32_t kl[4];
                    ldsm4(kl, k_tile(st, NPL - 1) + so);
                    mma16816(s[nb], qh[2 * kk2], kl[0], kl[1]);
                    if (2 * kk2 + 1 < KS) mma16816(s[nb], qh[2 * kk2 + 1], kl[2], kl[3]);
                }
            }
        }

        // ---- mask (diagonal and ragged last tile only)
        const int kbase = jt * BC;
        const int rA = row0 + g, rB = rA + 8;
        if ((causal && kbase + BC - 1 > row0 + off) || kbase + BC > Sk) {
#pragma unroll
            for (int nb = 0; nb < NB; nb++) {
#pragma unroll
                for (int e = 0; e < 4; e++) {
                    const int key = kbase + nb * 8 + 2 * t + (e & 1);
                    const int r = (e & 2) ? rB : rA;
                    if (key >= Sk || (causal && key > r + off)) s[nb][e] = -INFINITY;
                }
            }
        }

        // ---- online softmax (rows g, g+8; the 4 lanes of a quad share a row)
        float tm0 = -INFINITY, tm1 = -INFINITY;
#pragma unroll
        for (int nb = 0; nb < NB; nb++) {
            tm0 = fmaxf(tm0, fmaxf(s[nb][0], s[nb][1]));
            tm1 = fmaxf(tm1, fmaxf(s[nb][2], s[nb][3]));
        }
        tm0 = fmaxf(tm0, __shfl_xor_sync(0xffffffffu, tm0, 1));
        tm0 = fmaxf(tm0, __shfl_xor_sync(0xffffffffu, tm0, 2));
        tm1 = fmaxf(tm1, __shfl_xor_sync(0xffffffffu, tm1, 1));
        tm1 = fmaxf(tm1, __shfl_xor_sync(0xffffffffu, tm1, 2));
        const float n0 = fmaxf(m0, tm0), n1 = fmaxf(m1, tm1);
        const float sc0 = (n0 == -INFINITY) ? 1.f : expf((m0 - n0) * scale);
        const float sc1 = (n1 == -INFINITY) ? 1.f : expf((m1 - n1) * scale);
        const float b0 = (n0 == -INFINITY) ? 0.f : n0, b1 = (n1 == -INFINITY) ? 0.f : n1;
        m0 = n0; m1 = n1;
        l0 *= sc0; l1 *= sc1;
#pragma unroll
        for (int d = 0; d < DB; d++) { acc[d][0] *= sc0; acc[d][1] *= sc0; acc[d][2] *= sc1; acc[d][3] *= sc1; }
#pragma unroll
        for (int nb = 0; nb < NB; nb++) {
            s[nb][0] = expf((s[nb][0] - b0) * scale);
            s[nb][1] = expf((s[nb][1] - b0) * scale);
            s[nb][2] = expf((s[nb][2] - b1) * scale);
            s[nb][3] = expf((s[nb][3] - b1) * scale);
            l0 += s[nb][0] + s[nb][1];
            l1 += s[nb][2] + s[nb][3];
        }

        // ---- O += P V
#pragma unroll
        for (int kk = 0; kk < BC / 16; kk++) {
            uint32_t ph[4], pl[4];
            split2(s[2 * kk][0], s[2 * kk][1], ph[0], pl[0]);
            split2(s[2 * kk][2], s[2 * kk][3], ph[1], pl[1]);
            split2(s[2 * kk + 1][0], s[2 * kk + 1][1], ph[2], pl[2]);
            split2(s[2 * kk + 1][2], s[2 * kk + 1][3], ph[3], pl[3]);
            const int r = kk * 16 + (mat & 1) * 8 + mr;
#pragma unroll
            for (int dp = 0; dp < DB / 2; dp++) {
                const int c = 2 * dp + (mat >> 1);
                const uint32_t so = soff(r, c);
                uint32_t vf[4];
                ldsm4t(vf, v_tile(st, 0) + so);
                mma16816(acc[2 * dp], ph, vf[0], vf[1]);
                mma16816(acc[2 * dp], pl, vf[0], vf[1]);
                mma16816(acc[2 * dp + 1], ph, vf[2], vf[3]);
                mma16816(acc[2 * dp + 1], pl, vf[2], vf[3]);
                if (F32KV) {
                    uint32_t vl[4];
                    ldsm4t(vl, v_tile(st, NPL - 1) + so);
                    mma16816(acc[2 * dp], ph, vl[0], vl[1]);
                    mma16816(acc[2 * dp + 1], ph, vl[2], vl[3]);
                }
            }
        }
    }

    // ---- epilogue
    l0 += __shfl_xor_sync(0xffffffffu, l0, 1);
    l0 += __shfl_xor_sync(0xffffffffu, l0, 2);
    l1 += __shfl_xor_sync(0xffffffffu, l1, 1);
    l1 += __shfl_xor_sync(0xffffffffu, l1, 2);
    const float i0 = 1.0f / l0, i1 = 1.0f / l1;
    const int rA = row0 + g, rB = rA + 8;
#pragma unroll
    for (int d = 0; d < DB; d++) {
        const int c = d * 8 + 2 * t;
        if (rA < Sq) *(float2 *)(o + (((size_t)b * Sq + rA) * Hq + h) * D + c) = make_float2(acc[d][0] * i0, acc[d][1] * i0);
        if (rB < Sq) *(float2 *)(o + (((size_t)b * Sq + rB) * Hq + h) * D + c) = make_float2(acc[d][2] * i1, acc[d][3] * i1);
    }
}

// fp32 K/V -> fp16 planes (hi = fp16(x), lo = fp16(x - hi)): dst[0] K hi, [1] K lo, [2] V hi, [3] V lo, each [B][Sk][Hkv * D] dense
__global__ void k_split_kv(const float *__restrict__ k, const float *__restrict__ v, __half *__restrict__ dst, int Sk, long long row, long long kvb,
                           long long pstride, long long total2) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; // one float2 of K and of V
    if (i >= total2) return;
    const long long e = 2 * i, r = e / row, c = e % row, b = r / Sk, s = r % Sk; // row = Hkv * D (even)
    const size_t src = (size_t)b * kvb + (size_t)s * row + c;
    const float2 kk = *(const float2 *)(k + src), vv = *(const float2 *)(v + src);
    uint32_t kh, kl, vh, vl;
    split2(kk.x, kk.y, kh, kl);
    split2(vv.x, vv.y, vh, vl);
    uint32_t *d = (uint32_t *)(dst + e);
    d[0] = kh;
    d[pstride / 2] = kl;
    d[pstride] = vh;
    d[pstride / 2 * 3] = vl;
}

size_t attn_tc_f32_planes_bytes(int B, int Sq, int Sk, int Hkv, int D) {
    if (Sq < 16 || (D != 64 && D != 80 && D != 128)) return 0;
    const size_t ps = (((size_t)B * Sk * Hkv * D) + 7) & ~(size_t)7;
    return 4 * ps * sizeof(__half) + 256;
}

template <int D, int MODE>
static int launch_tc(const float *q, const void *k, const void *v, float *o, int B, int Sq, int Sk, int Hq, int Hkv, int causal,
                     long long kvb, long long pstride, cudaStream_t st) {
    constexpr int CH = D / 8, RCH = MODE == 2 ? CH + 1 : (CH <= 8) ? 8 : 16;
    // MODE 2: + 64 bytes -- for an odd number of k-steps (D = 80) the last ldmatrix of a row also names the chunk after the pad
    constexpr int smem = 64 * RCH * 16 * 2 * (MODE != 0 ? 2 : 1) * (MODE != 1 ? 2 : 1) + (MODE == 2 ? 64 : 0);
    static bool attr = false;
    if (!attr) {
        CU_TRY(cudaFuncSetAttribute(k_attn_prefill_tc<D, MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
        attr = true;
    }
    dim3 grid((Sq + 63) / 64, Hq, B);
    k_attn_prefill_tc<D, MODE><<<grid, 128, smem, st>>>(q, k, v, o, Sq, Sk, Hq, Hkv, causal, kvb, pstride);
    CU_LAUNCH_CHECK();
    return 0;
}

// Returns 1 when the shape is not for this kernel (caller falls through to the CUDA-core kernel), 0 on success.
int attn_prefill_tc_f16(const float *q, const void *k, const void *v, float *o, int B, int Sq, int Sk, int Hq, int Hkv, int D, int causal,
                        long long kvb, cudaStream_t st) {
    if (Sq < 16 || (kvb % 8) != 0) return 1;
    if (((uintptr_t)k | (uintptr_t)v) & 15) return 1;
    switch (D) {
    case 64: return launch_tc<64, 0>(q, k, v, o, B, Sq, Sk, Hq, Hkv, causal, kvb, 0, st);
    case 128: return launch_tc<128, 0>(q, k, v, o, B, Sq, Sk, Hq, Hkv, causal, kvb, 0, st);
    default: return 1;
    }
}
// ws / ws_bytes: room for the four planes (attn_tc_f32_planes_bytes) selects the pre-split kernel; less keeps the in-kernel split
int attn_prefill_tc_f32(const float *q, const void *k, const void *v, float *o, int B, int Sq, int Sk, int Hq, int Hkv, int D, int causal,
                        long long kvb, void *ws, size_t ws_bytes, cudaStream_t st) {
    if (Sq < 16 || (kvb % 4) != 0) return 1;
    if (((uintptr_t)k | (uintptr_t)v) & 15) return 1;
    const size_t need = attn_tc_f32_planes_bytes(B, Sq, Sk, Hkv, D);
    static const bool presplit_on = !(getenv("MLLMCU_ATTN_PRESPLIT") && atoi(getenv("MLLMCU_ATTN_PRESPLIT")) == 0);
    if (presplit_on && need && ws && ws_bytes >= need && !((uintptr_t)ws & 15) && (Hkv * D) % 8 == 0) {
        const long long row = (long long)Hkv * D, dense = (long long)Sk * row;
        const long long ps = (((long long)B * dense) + 7) & ~7LL, total2 = (long long)B * dense / 2;
        k_split_kv<<<(unsigned)((total2 + 255) / 256), 256, 0, st>>>((const float *)k, (const float *)v, (__half *)ws, Sk, row, kvb, ps, total2);
        CU_LAUNCH_CHECK();
        switch (D) {
        case 64: return launch_tc<64, 2>(q, ws, nullptr, o, B, Sq, Sk, Hq, Hkv, causal, dense, ps, st);
        case 80: return launch_tc<80, 2>(q, ws, nullptr, o, B, Sq, Sk, Hq, Hkv, causal, dense, ps, st);
        case 128: return launch_tc<128, 2>(q, ws, nullptr, o, B, Sq, Sk, Hq, Hkv, causal, dense, ps, st);
        }
    }
    switch (D) {
    case 64: return launch_tc<64, 1>(q, k, v, o, B, Sq, Sk, Hq, Hkv, causal, kvb, 0, st);
    case 80: return launch_tc<80, 1>(q, k, v, o, B, Sq, Sk, Hq, Hkv, causal, kvb, 0, st);
    case 128: return launch_tc<128, 1>(q, k, v, o, B, Sq, Sk, Hq, Hkv, causal, kvb, 0, st);
    default: return 1;
    }
}
