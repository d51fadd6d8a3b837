// decode_step.cu -- the whole single-token step of a LLaMA-family decoder as ONE persistent kernel.
//
// Why: a decode step is a chain of ~130 weight-streaming Linears (3.7 GB for the 7B) separated by tiny
// activation dependencies.  As separate kernels every Linear pays launch + prologue + pipeline ramp
// (4-5 us on a 2-8 us stream), which caps the chain at ~45 % of HBM bandwidth.  Here one CTA per SM
// stays resident for the whole step; every warp owns a ring of shared-memory slots fed by TMA bulk
// copies, and the ring runs AHEAD across Linear boundaries: weights depend on nothing, so while the
// grid waits at the barrier that publishes the previous Linear's outputs (and re-quantises the new
// activations), the copies for the next Linear are already landing.  HBM only idles when a dependency
// chain (barrier + prologue, or the attention phase) outlasts the ring (~200 KB per SM, ~4 us).
//
// Work list ("jobs", built by the host plan): per block  qkv -> [rope + KV append] -> [attention] ->
// o (+residual) -> gate|up (silu*mul) -> down (+residual), then final norm + lm_head.  Arithmetic is the same
// as gemv_tma.cu / decode.cu kernel for kernel (same helpers), so results equal the chained path bit for bit
// except the attention split count.
//
// Reference ops replaced: see gemv_tma.cu (Linear/RMSNorm/quantize/SiLU/add) and decode.cu (RoPE, KVCache,
// FlashAttention2).
#include "gemv_common.cuh"
#include <stdio.h>
#include <stdlib.h>

#define STEP_HDR 1536
#define STEP_PCH 4
#define STEP_MAX_STAGES 4

struct StepJob {
    const uint8_t *W[3];
    const float *bias[3];
    int N[3];
    int nW;
    int K;
    int epi;   // 0 plain, 1 + residual, 2 silu(gate) * up
    int post;  // 1: rope + KV append + attention of `layer` follow this job
    int layer;
    int rpw;   // rows per ring slot per matrix (1 or 2)
    const float *x;
    int ldx;
    const float *norm_w; // non-null: RMSNorm prologue
    float eps;
    const float *residual;
    int ldr;
    float *y;
    int ldy;
};

struct StepParams {
    const StepJob *jobs;
    int njobs;
    __half *const *kcache; // [layers]
    __half *const *vcache;
    int B, Hq, Hkv, D, rope_type, rot, tdim, limit, nsplit;
    const float *sin_t, *cos_t;
    long long kvb;
    float *qkv;
    int ldqkv;
    float *ao;
    int ldao;
    float *attn_ws;
    unsigned *tickets;
    const int *ctl;
    unsigned *gbar;
    int stages, slot_bytes, act_bytes;
    unsigned long long *prof; // optional [njobs][6] globaltimer stamps of CTA 0 (development aid)
};

static __device__ __forceinline__ unsigned ld_acquire_u32(const unsigned *p) {
    unsigned v;
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}

// all CTAs of the (co-resident) grid; `gen` counts expected arrivals and is uniform across threads
static __device__ __forceinline__ void grid_barrier(unsigned *ctr, unsigned &gen) {
    gen += gridDim.x;
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence();
        atomicAdd(ctr, 1u);
        const long long t0 = clock64();
        while (ld_acquire_u32(ctr) < gen) {
            if (clock64() - t0 > 4000000000LL) __trap(); // ~2 s: never hang the device on a logic error
        }
        __threadfence();
    }
    __syncthreads();
}

static __device__ __forceinline__ void mbar_wait_bounded(uint32_t bar, uint32_t phase) {
    for (int spin = 0;; spin++) {
        uint32_t ok;
        asm volatile(
            "{\n"
            ".reg .pred P1;\n"
            "mbarrier.try_wait.parity.shared::cta.b64 P1, [%1], %2;\n"
            "selp.u32 %0, 1, 0, P1;\n"
            "}"
            : "=r"(ok)
            : "r"(bar), "r"(phase)
            : "memory");
        if (ok) return;
        if (spin > (1 << 24)) __trap();
    }
}

static __device__ __forceinline__ unsigned long long gtimer() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    return t;
}
#define STEP_STAMP(k) do { if (P.prof && blockIdx.x == 0 && tid == 0) P.prof[(size_t)j * 6 + (k)] = gtimer(); } while (0)

static __device__ __forceinline__ void team_sync(int team, int nthreads) {
    asm volatile("bar.sync %0, %1;" ::"r"(team + 1), "r"(nthreads) : "memory");
}

// ---- one ring slot: NR weight rows against the M quantised activation rows in shared memory
template <int M, int NR>
static __device__ __forceinline__ void stage_compute(const uint8_t *wslot, int rowb, int nb, int lane, const uint8_t *act_lo,
                                                     const uint8_t *act_hi, const float2 *act_ds, float (&acc)[2][M]) {
    const int odd = lane & 1;
    const uint32_t sel = odd ? 0x7654u : 0x5432u;
#pragma unroll
    for (int j = 0; j < 2; j++)
#pragma unroll
        for (int m = 0; m < M; m++) acc[j][m] = 0.f;
#pragma unroll 2
    for (int b = lane; b < nb; b += 32) {
        uint32_t lo[NR][4], hi[NR][4];
        float dw[NR];
#pragma unroll
        for (int j = 0; j < NR; j++) load_block(wslot + (size_t)j * rowb, b, odd, sel, lo[j], hi[j], dw[j]);
#pragma unroll
        for (int m = 0; m < M; m++) {
            const int4 al = *(const int4 *)(act_lo + ((size_t)m * nb + b) * 16);
            const int4 ah = *(const int4 *)(act_hi + ((size_t)m * nb + b) * 16);
            const float2 ds = act_ds[(size_t)m * nb + b];
            const int asum = __float_as_int(ds.y);
#pragma unroll
            for (int j = 0; j < NR; j++) acc[j][m] = fmaf(dw[j] * ds.x, (float)block_idot(lo[j], hi[j], al, ah, asum), acc[j][m]);
        }
    }
}

// ---- (RMSNorm) + Q8_0 quantisation of the M input rows of a job into shared memory (same arithmetic as gemv_tma.cu);
// inputs were written by other CTAs earlier in this launch: read them past L1 (ld.global.cg)
template <int M>
static __device__ void step_prologue(const StepJob &jb, uint8_t *act_lo, uint8_t *act_hi, float2 *act_ds, float *red, float *s_rms) {
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nwarps = blockDim.x >> 5;
    const int K = jb.K, nb = K / QK;
    const bool pro = jb.norm_w != nullptr;
    const int g = lane >> 3, sub = lane & 7;
    const int bstep = nwarps * 4;
    const float4 z4 = make_float4(0.f, 0.f, 0.f, 0.f);
    for (int m = 0; m < M; m++) {
        const float *xr = jb.x + (size_t)m * jb.ldx;
        uint8_t *lo_m = act_lo + (size_t)m * nb * 16, *hi_m = act_hi + (size_t)m * nb * 16;
        float2 *ds_m = act_ds + (size_t)m * nb;
        const bool single = nb <= bstep * STEP_PCH;
        float4 v[STEP_PCH], wv[STEP_PCH];
        float rms = 1.0f;
        if (pro) {
            float s = 0.f, c = 0.f;
            if (single) {
#pragma unroll
                for (int j = 0; j < STEP_PCH; j++) {
                    const int b = warp * 4 + g + j * bstep;
                    v[j] = (b < nb) ? __ldcg((const float4 *)(xr + (size_t)b * QK + sub * 4)) : z4;
                    wv[j] = (b < nb) ? *(const float4 *)(jb.norm_w + (size_t)b * QK + sub * 4) : z4;
                }
#pragma unroll
                for (int j = 0; j < STEP_PCH; j++) { sq_acc(v[j].x, s, c); sq_acc(v[j].y, s, c); sq_acc(v[j].z, s, c); sq_acc(v[j].w, s, c); }
            } else {
                for (int d = tid * 4; d < K; d += blockDim.x * 4) {
                    const float4 t = __ldcg((const float4 *)(xr + d));
                    sq_acc(t.x, s, c); sq_acc(t.y, s, c); sq_acc(t.z, s, c); sq_acc(t.w, s, c);
                }
            }
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
                const float s2 = __shfl_xor_sync(0xffffffffu, s, o), c2 = __shfl_xor_sync(0xffffffffu, c, o);
                ds_add(s, c, s2, c2);
            }
            if (lane == 0) { red[warp * 2] = s; red[warp * 2 + 1] = c; }
            __syncthreads();
            if (tid == 0) {
                double t = 0.0;
                for (int w = 0; w < nwarps; w++) t += (double)red[w * 2] + (double)red[w * 2 + 1];
                const float mean = (float)(t / (double)K);
                s_rms[m] = __fdiv_rn(1.0f, __fsqrt_rn(__fadd_rn(mean, jb.eps)));
            }
            __syncthreads();
            rms = s_rms[m];
        }
        for (int base = warp * 4; base < nb; base += bstep * STEP_PCH) { // warp-uniform trip count (shuffles inside)
            if (!(pro && single)) {
#pragma unroll
                for (int j = 0; j < STEP_PCH; j++) {
                    const int b = base + g + j * bstep;
                    v[j] = (b < nb) ? __ldcg((const float4 *)(xr + (size_t)b * QK + sub * 4)) : z4;
                    if (pro) wv[j] = (b < nb) ? *(const float4 *)(jb.norm_w + (size_t)b * QK + sub * 4) : z4;
                }
            }
#pragma unroll
            for (int j = 0; j < STEP_PCH; j++) {
                if (base + j * bstep >= nb) break; // warp-uniform
                const int b = base + g + j * bstep;
                const bool ok = b < nb;
                float4 t = v[j];
                if (pro) {
                    const float4 w = wv[j];
                    t.x = __fmul_rn(__fmul_rn(t.x, rms), w.x); t.y = __fmul_rn(__fmul_rn(t.y, rms), w.y);
                    t.z = __fmul_rn(__fmul_rn(t.z, rms), w.z); t.w = __fmul_rn(__fmul_rn(t.w, rms), w.w);
                }
                quant_store(t, ok, b, sub, lo_m, hi_m, ds_m);
            }
        }
    }
    __syncthreads();
}

// ---- RoPE(q, k) + KV append for one layer, grid-wide (arithmetic of k_decode_rope_kv)
static __device__ void step_rope_kv(const StepParams &P, int layer) {
    const int D = P.D, Hq = P.Hq, Hkv = P.Hkv, half_d = D / 2, HT = Hq + 2 * Hkv, hr = P.rot / 2;
    const int pos = P.ctl[0];
    if (pos < 0 || pos >= P.limit) return;
    __half *kc = P.kcache[layer], *vc = P.vcache[layer];
    const long long total = (long long)P.B * HT * half_d;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        const int p = (int)(i % half_d);
        const int h = (int)((i / half_d) % HT);
        const int b = (int)(i / ((long long)half_d * HT));
        float *xi = P.qkv + (size_t)b * P.ldqkv + (size_t)h * D;
        if (h >= Hq + Hkv) {
            __half *dst = vc + (size_t)b * P.kvb + (size_t)pos * Hkv * D + (size_t)(h - Hq - Hkv) * D;
            const float2 v = __ldcg((const float2 *)(xi + 2 * p));
            *(__half2 *)(dst + 2 * p) = __floats2half2_rn(v.x, v.y);
            continue;
        }
        int d0, d1;
        float v0, v1;
        if (p < hr) {
            if (P.rope_type == 2) { d0 = 2 * p; d1 = 2 * p + 1; }
            else { d0 = p; d1 = p + hr; }
            const float a = __ldcg(xi + d0), bb = __ldcg(xi + d1);
            const float sn = P.sin_t[(size_t)pos * P.tdim + d0], cs = P.cos_t[(size_t)pos * P.tdim + d0];
            v0 = rope_lo(a, bb, sn, cs);
            v1 = rope_hi(a, bb, sn, cs);
        } else {
            d0 = P.rot + 2 * (p - hr); d1 = d0 + 1;
            v0 = __ldcg(xi + d0); v1 = __ldcg(xi + d1);
        }
        if (h < Hq) {
            xi[d0] = v0; xi[d1] = v1;
        } else {
            __half *dst = kc + (size_t)b * P.kvb + (size_t)pos * Hkv * D + (size_t)(h - Hq) * D;
            dst[d0] = __float2half_rn(v0);
            dst[d1] = __float2half_rn(v1);
        }
    }
}

static __device__ __forceinline__ void h8f(const uint4 &u, float *f) {
    const __half2 *h = (const __half2 *)&u;
#pragma unroll
    for (int i = 0; i < 4; i++) {
        const float2 t = __half22float2(h[i]);
        f[2 * i] = t.x; f[2 * i + 1] = t.y;
    }
}

// ---- decode attention of one layer: a team of 4 warps per (batch row, q head, key split); arithmetic of
// k_attn_decode (decode.cu).  `scratch` is the team's [4][D+2] floats + 1 flag in shared memory.
template <int D>
static __device__ void step_attention(const StepParams &P, int layer, float *scratch_all) {
    constexpr int LPK = D / 16, KPW = 32 / LPK;
    const int nteams = blockDim.x >> 7;
    const int team = threadIdx.x >> 7, tid = threadIdx.x & 127, lane = tid & 31, warp = tid >> 5;
    if (team >= nteams) return; // warps beyond a multiple of 4 idle here
    const int item = blockIdx.x * nteams + team;
    const int nsplit = P.nsplit, Hq = P.Hq, Hkv = P.Hkv;
    if (item >= P.B * Hq * nsplit) return; // whole team leaves together
    float(*sm)[D + 2] = (float(*)[D + 2])(scratch_all + (size_t)team * (4 * (D + 2) + 4));
    int *s_last = (int *)(scratch_all + (size_t)team * (4 * (D + 2) + 4) + 4 * (D + 2));
    const int sp = item % nsplit, h = (item / nsplit) % Hq, b = item / (nsplit * Hq);
    const int kvh = h / (Hq / Hkv);
    const int Sk = P.ctl[0] + 1;
    const int chunk = (Sk + nsplit - 1) / nsplit;
    const int c_begin = sp * chunk, c_end = min(Sk, c_begin + chunk);
    const int g = lane / LPK, sub = lane % LPK;
    const float scale = __fdiv_rn(1.0f, __fsqrt_rn((float)D));
    const __half *k = P.kcache[layer], *v = P.vcache[layer];

    float qv[16], acc[16];
    {
        const float4 *qp = (const float4 *)(P.qkv + (size_t)b * P.ldqkv + (size_t)h * D + sub * 16);
#pragma unroll
        for (int i = 0; i < 4; i++) { const float4 t = __ldcg(qp + i); qv[4 * i] = t.x; qv[4 * i + 1] = t.y; qv[4 * i + 2] = t.z; qv[4 * i + 3] = t.w; }
    }
#pragma unroll
    for (int i = 0; i < 16; i++) acc[i] = 0.f;
    float m = -INFINITY, l = 0.f;
    const size_t rs = (size_t)Hkv * D;
    const __half *kb = k + (size_t)b * P.kvb + (size_t)kvh * D + sub * 16;
    const __half *vb = v + (size_t)b * P.kvb + (size_t)kvh * D + sub * 16;
#pragma unroll 2
    for (int c0 = c_begin + warp * KPW; c0 < c_end; c0 += 4 * KPW) {
        const int c = c0 + g;
        const bool ok = c < c_end;
        const size_t off = (size_t)(ok ? c : c_begin) * rs;
        const uint4 k0 = __ldcg((const uint4 *)(kb + off)), k1 = __ldcg((const uint4 *)(kb + off + 8));
        const uint4 v0 = __ldcg((const uint4 *)(vb + off)), v1 = __ldcg((const uint4 *)(vb + off + 8));
        float kf[16], vf[16];
        h8f(k0, kf); h8f(k1, kf + 8);
        float s = 0.f;
#pragma unroll
        for (int i = 0; i < 16; i++) s = fmaf(qv[i], kf[i], s);
#pragma unroll
        for (int x = 1; x < LPK; x <<= 1) s += __shfl_xor_sync(0xffffffffu, s, x);
        if (ok) {
            h8f(v0, vf); h8f(v1, vf + 8);
            const float m_new = fmaxf(m, s);
            const float sc = expf((m - m_new) * scale);
            const float p = expf((s - m_new) * scale);
            l = fmaf(l, sc, p);
#pragma unroll
            for (int i = 0; i < 16; i++) acc[i] = fmaf(p, vf[i], acc[i] * sc);
            m = m_new;
        }
    }
#pragma unroll
    for (int x = LPK; x < 32; x <<= 1) {
        const float m2 = __shfl_xor_sync(0xffffffffu, m, x), l2 = __shfl_xor_sync(0xffffffffu, l, x);
        const float mn = fmaxf(m, m2);
        const float f1 = (m == -INFINITY) ? 0.f : expf((m - mn) * scale);
        const float f2 = (m2 == -INFINITY) ? 0.f : expf((m2 - mn) * scale);
        l = fmaf(l, f1, l2 * f2);
#pragma unroll
        for (int i = 0; i < 16; i++) {
            const float a2 = __shfl_xor_sync(0xffffffffu, acc[i], x);
            acc[i] = fmaf(acc[i], f1, a2 * f2);
        }
        m = mn;
    }
    if (g == 0) {
        if (sub == 0) { sm[warp][0] = m; sm[warp][1] = l; }
#pragma unroll
        for (int i = 0; i < 16; i++) sm[warp][2 + sub * 16 + i] = acc[i];
    }
    team_sync(team, 128);
    float mt = -INFINITY, lt = 0.f, at = 0.f;
    if (tid < D) {
#pragma unroll
        for (int w = 0; w < 4; w++) mt = fmaxf(mt, sm[w][0]);
#pragma unroll
        for (int w = 0; w < 4; w++) {
            const float mw = sm[w][0];
            const float f = (mw == -INFINITY) ? 0.f : expf((mw - mt) * scale);
            lt = fmaf(sm[w][1], f, lt);
            at = fmaf(sm[w][2 + tid], f, at);
        }
    }
    float *op = P.ao + (size_t)b * P.ldao + (size_t)h * D;
    if (nsplit == 1) {
        if (tid < D) op[tid] = at * (1.0f / lt);
        return;
    }
    const size_t row = (size_t)b * Hq + h;
    float *wp = P.attn_ws + (row * nsplit + sp) * (D + 2);
    if (tid < D) {
        if (tid == 0) { wp[0] = mt; wp[1] = lt; }
        wp[2 + tid] = at;
    }
    __threadfence();
    team_sync(team, 128);
    if (tid == 0) {
        const unsigned t = atomicAdd(&P.tickets[row], 1u);
        *s_last = (t == (unsigned)nsplit - 1u);
        if (*s_last) P.tickets[row] = 0;
    }
    team_sync(team, 128);
    if (!*s_last) return;
    __threadfence();
    if (tid < D) {
        const float *wr = P.attn_ws + row * nsplit * (D + 2);
        float mm = -INFINITY;
        for (int s = 0; s < nsplit; s++) mm = fmaxf(mm, __ldcg(wr + (size_t)s * (D + 2)));
        float ll = 0.f, aa = 0.f;
        for (int s = 0; s < nsplit; s++) {
            const float ms = __ldcg(wr + (size_t)s * (D + 2));
            if (ms == -INFINITY) continue;
            const float f = expf((ms - mm) * scale);
            ll = fmaf(__ldcg(wr + (size_t)s * (D + 2) + 1), f, ll);
            aa = fmaf(__ldcg(wr + (size_t)s * (D + 2) + 2 + tid), f, aa);
        }
        op[tid] = aa * (1.0f / ll);
    }
}

template <int M>
__global__ void __launch_bounds__(512, 1) k_decode_step(const StepParams P) {
    extern __shared__ __align__(128) uint8_t smem[];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nwarps = blockDim.x >> 5;
    const int S = P.stages;
    uint64_t *bars = (uint64_t *)smem;              // [nwarps][S]
    float *red = (float *)(smem + 1024);            // [nwarps][2]
    float *s_rms = (float *)(smem + 1024 + 256);    // [M]
    uint8_t *act = smem + STEP_HDR;                 // activation image of the current job / attention scratch
    uint8_t *wmine = smem + STEP_HDR + P.act_bytes + (size_t)warp * S * P.slot_bytes;
    const int g0 = blockIdx.x * nwarps + warp, gstride = gridDim.x * nwarps;

    auto groups_of = [&](const StepJob &jb) { // row groups of this warp in a job
        int NT = jb.N[0];
        if (jb.epi != 2) for (int i = 1; i < jb.nW; i++) NT += jb.N[i];
        const int G = (NT + jb.rpw - 1) / jb.rpw;
        return g0 < G ? (G - 1 - g0) / gstride + 1 : 0;
    };
    // ---- producer state (lane 0): next (job, group) to fetch and the running slot counter
    int pj = 0, pi = 0, pn = 0;
    unsigned pcount = 0;
    if (lane == 0 && P.njobs > 0) pn = groups_of(P.jobs[0]);
    auto produce = [&]() { // lane 0: start the copy of the next row group, if any is left in the step
        while (pj < P.njobs && pi >= pn) {
            pj++;
            pi = 0;
            if (pj < P.njobs) pn = groups_of(P.jobs[pj]);
        }
        if (pj >= P.njobs) return;
        const StepJob &jb = P.jobs[pj];
        const int rowb = jb.K / QK * 18;
        const int row0 = (g0 + pi * gstride) * jb.rpw;
        int mi = 0, local = row0;
        if (jb.epi != 2) while (mi < jb.nW - 1 && local >= jb.N[mi]) { local -= jb.N[mi]; mi++; }
        const int nrows = min(jb.rpw, jb.N[mi] - local);
        const uint32_t bytes = (uint32_t)nrows * rowb;
        const int slot = (int)(pcount % (unsigned)S);
        const uint32_t bar = smem_u32(&bars[warp * S + slot]);
        uint8_t *dst = wmine + (size_t)slot * P.slot_bytes;
        mbar_expect_tx(bar, jb.epi == 2 ? 2 * bytes : bytes);
        bulk_g2s(smem_u32(dst), jb.W[mi] + (size_t)local * rowb, bytes, bar);
        if (jb.epi == 2) bulk_g2s(smem_u32(dst + (size_t)jb.rpw * rowb), jb.W[1] + (size_t)local * rowb, bytes, bar);
        pi++;
        pcount++;
    };
    if (lane == 0) {
        for (int s = 0; s < S; s++) mbar_init(smem_u32(&bars[warp * S + s]), 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        for (int s = 0; s < S; s++) produce();
    }
    unsigned ccount = 0, gen = 0;
    for (int j = 0; j < P.njobs; j++) {
        const StepJob &jb = P.jobs[j];
        STEP_STAMP(0);
        if (j > 0) grid_barrier(P.gbar, gen); // inputs of this job are complete and visible
        STEP_STAMP(1);
        const int nb = jb.K / QK, rowb = nb * 18;
        uint8_t *act_lo = act, *act_hi = act + (size_t)M * nb * 16;
        float2 *act_ds = (float2 *)(act_hi + (size_t)M * nb * 16);
        step_prologue<M>(jb, act_lo, act_hi, act_ds, red, s_rms);
        STEP_STAMP(2);
        const int n_my = groups_of(jb);
        const int nr = jb.rpw * (jb.epi == 2 ? 2 : 1);
        for (int i = 0; i < n_my; i++, ccount++) {
            const int slot = (int)(ccount % (unsigned)S);
            mbar_wait_bounded(smem_u32(&bars[warp * S + slot]), (ccount / (unsigned)S) & 1u);
            const uint8_t *wslot = wmine + (size_t)slot * P.slot_bytes;
            float acc[2][M];
            if (nr == 2) stage_compute<M, 2>(wslot, rowb, nb, lane, act_lo, act_hi, act_ds, acc);
            else stage_compute<M, 1>(wslot, rowb, nb, lane, act_lo, act_hi, act_ds, acc);
            __syncwarp();
            if (lane == 0) produce(); // the slot is free again: fetch ahead, possibly for a later job
            const int row0 = (g0 + i * gstride) * jb.rpw;
            int mi = 0, local = row0;
            if (jb.epi != 2) while (mi < jb.nW - 1 && local >= jb.N[mi]) { local -= jb.N[mi]; mi++; }
            const int nrows = min(jb.rpw, jb.N[mi] - local);
            for (int r = 0; r < nrows; r++) {
#pragma unroll
                for (int m = 0; m < M; m++) {
                    float v = warp_sum(r == 0 ? acc[0][m] : acc[1][m]);
                    float v2 = (jb.epi == 2) ? warp_sum(acc[1][m]) : 0.f; // gate/up jobs run one row pair per slot
                    if (lane == 0) {
                        const int nloc = local + r, ncol = row0 + r;
                        if (jb.bias[mi]) v = v + jb.bias[mi][nloc];
                        if (jb.epi == 1) v = __fadd_rn(v, __ldcg(jb.residual + (size_t)m * jb.ldr + ncol));
                        if (jb.epi == 2) {
                            if (jb.bias[1]) v2 = v2 + jb.bias[1][nloc];
                            v = __fmul_rn(g_silu(v), v2);
                        }
                        jb.y[(size_t)m * jb.ldy + ncol] = v;
                    }
                }
            }
        }
        STEP_STAMP(3);
        if (jb.post) {
            grid_barrier(P.gbar, gen); // q|k|v complete
            step_rope_kv(P, jb.layer);
            STEP_STAMP(4);
            grid_barrier(P.gbar, gen); // rotated q and the new cache row visible
            STEP_STAMP(5);
            if (P.D == 128) step_attention<128>(P, jb.layer, (float *)act);
            else if (P.D == 64) step_attention<64>(P, jb.layer, (float *)act);
            else step_attention<32>(P, jb.layer, (float *)act);
        }
    }
}

// ------------------------------------------------------------------ host side
#include "mllmcu.h"
#include <vector>

struct StepPlan {
    StepParams P = {};
    int M = 0, grid = 0, threads = 0;
    size_t smem = 0;
    void *dev_blob = nullptr; // jobs | kcache ptrs | vcache ptrs | barrier | attention workspace + tickets
    unsigned long long *prof = nullptr;
    int launches = 0;
};

template <int M>
static int step_launch_m(StepPlan *pl, cudaStream_t st) {
    static size_t configured = 0;
    if (pl->smem > configured) {
        CU_TRY(cudaFuncSetAttribute(k_decode_step<M>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)pl->smem));
        configured = pl->smem;
    }
    CU_TRY(cudaMemsetAsync(pl->P.gbar, 0, sizeof(unsigned), st));
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(pl->grid);
    cfg.blockDim = dim3(pl->threads);
    cfg.dynamicSmemBytes = pl->smem;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeCooperative; // every CTA must be resident: the kernel has grid-wide barriers
    attr[0].val.cooperative = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    CU_TRY(cudaLaunchKernelEx(&cfg, k_decode_step<M>, pl->P));
    if (pl->prof && ++pl->launches == 6) { // development aid (MLLMCU_STEP_PROF=1, run without graph capture)
        CU_TRY(cudaStreamSynchronize(st));
        std::vector<unsigned long long> h((size_t)pl->P.njobs * 6);
        CU_TRY(cudaMemcpy(h.data(), pl->prof, h.size() * 8, cudaMemcpyDeviceToHost));
        const unsigned long long t0 = h[0];
        for (int j = 0; j < pl->P.njobs && j < 12; j++) {
            fprintf(stderr, "[step prof] job %3d:", j);
            for (int k = 0; k < 6; k++) fprintf(stderr, " %8.2f", h[(size_t)j * 6 + k] ? (double)(h[(size_t)j * 6 + k] - t0) / 1000.0 : -1.0);
            fprintf(stderr, "  us (start, barrier, prologue, stream, rope, barrier2)\n");
        }
        const int lj = pl->P.njobs - 1;
        fprintf(stderr, "[step prof] last job %d start %.2f stream done %.2f us\n", lj, (double)(h[(size_t)lj * 6] - t0) / 1000.0,
                (double)(h[(size_t)lj * 6 + 3] - t0) / 1000.0);
    }
    return 0;
}

MLLMCU_API int mllmcu_decode_step_launch(void *plan, void *stream) {
    StepPlan *pl = (StepPlan *)plan;
    if (!pl) return -1;
    cudaStream_t st = (cudaStream_t)stream;
    switch (pl->M) {
    case 1: return step_launch_m<1>(pl, st);
    case 2: return step_launch_m<2>(pl, st);
    case 3: return step_launch_m<3>(pl, st);
    case 4: return step_launch_m<4>(pl, st);
    default: return -2;
    }
}

MLLMCU_API int mllmcu_decode_step_destroy(void *plan) {
    StepPlan *pl = (StepPlan *)plan;
    if (!pl) return 0;
    if (pl->dev_blob) cudaFree(pl->dev_blob);
    if (pl->prof) cudaFree(pl->prof);
    delete pl;
    return 0;
}

MLLMCU_API int mllmcu_decode_step_create(const mllmcu_step_desc *d, const mllmcu_step_job *jobs, int njobs, void **plan_out) {
    if (!d || !jobs || njobs <= 0 || !plan_out) return -1;
    if (d->B < 1 || d->B > 4) return -2;
    if (d->D != 32 && d->D != 64 && d->D != 128) return -2;
    if (d->Hq % d->Hkv || (d->ldqkv & 3) || (d->kv_bstride & 7) || ((d->Hkv * d->D) & 7)) return -2;
    if ((d->rope_type != 2 && d->rope_type != 4) || d->rot_dim % 2 || d->rot_dim > d->D || d->rot_dim > d->tdim) return -1;
    const int M = d->B;
    int dev = 0, sms = 0;
    CU_TRY(cudaGetDevice(&dev));
    CU_TRY(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    // ---- shared-memory budget: activation image of the widest input (or the attention scratch), then the ring
    int maxK = 0;
    for (int j = 0; j < njobs; j++) {
        const mllmcu_step_job &jb = jobs[j];
        if (jb.K <= 0 || jb.K % 128 || jb.nW < 1 || jb.nW > 3 || jb.epi < 0 || jb.epi > 2) return -2;
        if (jb.epi == 2 && (jb.nW != 2 || jb.N[0] != jb.N[1])) return -1;
        if (jb.epi == 1 && !jb.residual) return -1;
        if (((uintptr_t)jb.x & 15) || (jb.ldx & 3) || (jb.norm_w && ((uintptr_t)jb.norm_w & 15))) return -2;
        for (int i = 0; i < jb.nW; i++) if (((uintptr_t)jb.W[i] & 15) || jb.N[i] <= 0) return -2;
        if (jb.K > maxK) maxK = jb.K;
    }
    const char *e = getenv("MLLMCU_STEP_WARPS");
    const int want_warps = (e && e[0]) ? atoi(e) : 0;
    e = getenv("MLLMCU_STEP_STAGES");
    const int want_stages = (e && e[0]) ? atoi(e) : 0;
    const size_t budget = 220 * 1024;
    StepPlan *pl = nullptr;
    std::vector<StepJob> hj(njobs);
    for (int nwarps : {16, 12, 8}) {
        if (want_warps && nwarps != want_warps) continue;
        const int nteams = nwarps / 4;
        size_t act = (size_t)M * (maxK / QK) * 40;
        const size_t scratch = (size_t)nteams * (4 * (d->D + 2) + 4) * sizeof(float);
        if (scratch > act) act = scratch;
        act = (act + 127) & ~(size_t)127;
        for (int stages : {2, 3, 4}) {
            if (want_stages && stages != want_stages) continue;
            if (STEP_HDR + act >= budget) continue;
            size_t slot = (budget - STEP_HDR - act) / ((size_t)nwarps * stages);
            slot &= ~(size_t)127;
            // rows per slot for every job under this slot size
            bool ok = slot >= 128;
            for (int j = 0; ok && j < njobs; j++) {
                const mllmcu_step_job &jb = jobs[j];
                const size_t rowb = (size_t)jb.K / QK * 18;
                bool even = true;
                for (int i = 0; i < (jb.epi == 2 ? 1 : jb.nW); i++) if (jb.N[i] % 2) even = false;
                int rpw;
                if (jb.epi == 2) rpw = 1;
                else rpw = (M <= 2 && even && 2 * rowb <= slot) ? 2 : 1;
                if ((rowb & 15) && !(rpw == 2)) ok = false; // rows of 8 mod 16 bytes need aligned pairs
                if ((size_t)rpw * (jb.epi == 2 ? 2 : 1) * rowb > slot) ok = false;
                StepJob &o = hj[j];
                for (int i = 0; i < 3; i++) { o.W[i] = (const uint8_t *)jb.W[i]; o.bias[i] = jb.bias[i]; o.N[i] = jb.N[i]; }
                o.nW = jb.nW; o.K = jb.K; o.epi = jb.epi; o.post = jb.post; o.layer = jb.layer; o.rpw = rpw;
                o.x = jb.x; o.ldx = jb.ldx; o.norm_w = jb.norm_w; o.eps = jb.eps; o.residual = jb.residual; o.ldr = jb.ldr;
                o.y = jb.y; o.ldy = jb.ldy;
                if (jb.post && (jb.layer < 0 || jb.layer >= d->n_layers)) return -1;
            }
            if (!ok) continue;
            pl = new StepPlan();
            pl->M = M;
            pl->grid = sms;
            pl->threads = nwarps * 32;
            pl->P.stages = stages;
            pl->P.slot_bytes = (int)slot;
            pl->P.act_bytes = (int)act;
            pl->smem = STEP_HDR + act + (size_t)nwarps * stages * slot;
            break;
        }
        if (pl) break;
    }
    if (!pl) return -2;
    const int nteams = pl->threads / 128;
    int nsplit = (pl->grid * nteams) / (d->B * d->Hq);
    if (nsplit < 1) { delete pl; return -2; } // more (row, head) pairs than attention teams
    if (nsplit > 32) nsplit = 32;
    // ---- device tables
    const size_t o_jobs = 0, o_kc = (sizeof(StepJob) * njobs + 255) & ~(size_t)255, o_vc = o_kc + ((sizeof(void *) * d->n_layers + 255) & ~(size_t)255);
    const size_t o_bar = o_vc + ((sizeof(void *) * d->n_layers + 255) & ~(size_t)255), o_ws = o_bar + 256;
    const size_t rows = (size_t)d->B * d->Hq;
    const size_t ws_bytes = rows * nsplit * (d->D + 2) * sizeof(float), o_tk = o_ws + ((ws_bytes + 255) & ~(size_t)255);
    const size_t total = o_tk + rows * sizeof(unsigned) + 256;
    cudaError_t rc = cudaMalloc(&pl->dev_blob, total);
    if (rc != cudaSuccess) { delete pl; return (int)rc; }
    char *blob = (char *)pl->dev_blob;
    rc = cudaMemset(blob, 0, total);
    if (rc == cudaSuccess) rc = cudaMemcpy(blob + o_jobs, hj.data(), sizeof(StepJob) * njobs, cudaMemcpyHostToDevice);
    if (rc == cudaSuccess) rc = cudaMemcpy(blob + o_kc, d->kcache, sizeof(void *) * d->n_layers, cudaMemcpyHostToDevice);
    if (rc == cudaSuccess) rc = cudaMemcpy(blob + o_vc, d->vcache, sizeof(void *) * d->n_layers, cudaMemcpyHostToDevice);
    if (rc != cudaSuccess) { mllmcu_decode_step_destroy(pl); return (int)rc; }
    StepParams &P = pl->P;
    P.jobs = (const StepJob *)(blob + o_jobs);
    P.njobs = njobs;
    P.kcache = (__half *const *)(blob + o_kc);
    P.vcache = (__half *const *)(blob + o_vc);
    P.B = d->B; P.Hq = d->Hq; P.Hkv = d->Hkv; P.D = d->D; P.rope_type = d->rope_type; P.rot = d->rot_dim; P.tdim = d->tdim;
    P.limit = d->cache_limit; P.nsplit = nsplit; P.sin_t = d->sin_t; P.cos_t = d->cos_t; P.kvb = d->kv_bstride;
    P.qkv = d->qkv; P.ldqkv = d->ldqkv; P.ao = d->ao; P.ldao = d->ldao;
    P.attn_ws = (float *)(blob + o_ws);
    P.tickets = (unsigned *)(blob + o_tk);
    P.ctl = d->ctl;
    P.gbar = (unsigned *)(blob + o_bar);
    e = getenv("MLLMCU_STEP_PROF");
    if (e && e[0] == '1') {
        if (cudaMalloc(&pl->prof, (size_t)njobs * 6 * 8) == cudaSuccess) cudaMemset(pl->prof, 0, (size_t)njobs * 6 * 8);
        P.prof = pl->prof;
    }
    *plan_out = pl;
    return 0;
}
