// decode.cu -- the single-token (S == 1) step of the decoder, shaped for CUDA-graph replay:
// every per-step quantity (position, KV length, token id) lives in device memory, so the
// captured graph is identical for every token.
//
//   ctl[0]  position of the token being consumed == KV-cache length before its append.
//           k_decode_embed increments it at the start of each step; the host sets it to
//           (cache length - 1) whenever its shadow copy is stale (after prefill / clear).
//
// Kernels (reference ops they replace, fused):
//   k_decode_embed    CPUEmbedding Q4_0 row gather (op/CPUEmbedding.cpp:66-81)
//   k_decode_rope_kv  q_rope, k_rope (op/CPURoPE.cpp:200-372) + k_cache/v_cache append in the
//                     cache dtype fp16 (op/CPUKVCache.cpp:137-247) from the fused q|k|v projection
//   k_attn_decode     flash_attention_2_forward decode branch (compute/FlashAttention2.hpp,
//                     fp32 q, fp16 K/V) split over keys, partials merged by the last CTA to finish
//   k_argmax (elementwise.cu) with token log: greedy pick (Generate.cpp:46-53)
#include "common.cuh"
#include <stdlib.h>

static __device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
static __device__ __forceinline__ void pdl_launch() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }

// ------------------------------------------------------------------ embedding + step counter
__global__ void k_decode_embed(const uint8_t *__restrict__ table, const float *__restrict__ ids, float *__restrict__ out, int n_ids,
                               int hidden, int *ctl) {
    const int nb = hidden / QK;
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i == 0 && ctl) ctl[0] = ctl[0] + 1;
    if (i >= (long long)n_ids * nb * 16) return;
    const int j = (int)(i & 15);
    const int b = (int)((i >> 4) % nb);
    const int t = (int)((i >> 4) / nb);
    const float idf = ids[t];
    float *o = out + (size_t)t * hidden + (size_t)b * QK;
    if (!(idf >= 0)) { o[j] = 0.f; o[j + 16] = 0.f; return; } // padding id: zero row (CPUEmbedding.cpp:52-60)
    const uint8_t *blk = table + ((size_t)(int)idf * nb + b) * 18;
    const float d = __half2float(*(const __half *)blk);
    const uint8_t q = blk[2 + j];
    o[j] = __fmul_rn((float)((int)(q & 0x0F) - 8), d);
    o[j + 16] = __fmul_rn((float)((int)(q >> 4) - 8), d);
}
MLLMCU_API int mllmcu_decode_embed(const void *table, const float *ids, float *out, int n_ids, int hidden, int *ctl, void *stream) {
    if (n_ids <= 0 || hidden % QK) return -1;
    const long long tot = (long long)n_ids * (hidden / QK) * 16;
    k_decode_embed<<<(unsigned)((tot + 255) / 256), 256, 0, (cudaStream_t)stream>>>((const uint8_t *)table, ids, out, n_ids, hidden, ctl);
    CU_LAUNCH_CHECK();
    return 0;
}

// ------------------------------------------------------------------ RoPE(q,k) + KV append
// qkv: [B][ld] fp32 rows holding q (Hq*D) | k (Hkv*D) | v (Hkv*D).  q is rotated in place; k is
// rotated and stored fp16 at cache row `pos`; v is converted.  One thread per element pair.
__global__ void k_decode_rope_kv(int type, float *__restrict__ qkv, int ld, int B, int Hq, int Hkv, int D, int rot,
                                 const float *__restrict__ sin_t, const float *__restrict__ cos_t, int tdim, const int *__restrict__ ctl,
                                 __half *__restrict__ kc, __half *__restrict__ vc, long long kvb, int limit) {
    pdl_wait();
    pdl_launch();
    const int half_d = D / 2, HT = Hq + 2 * Hkv;
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (long long)B * HT * half_d) return;
    const int p = (int)(i % half_d);
    const int h = (int)((i / half_d) % HT);
    const int b = (int)(i / ((long long)half_d * HT));
    const int pos = ctl[0];
    if (pos < 0 || pos >= limit) return; // the host refuses the step before launching; never write outside the cache
    float *xi = qkv + (size_t)b * ld + (size_t)h * D;
    if (h >= Hq + Hkv) { // v: plain conversion
        const int hv = h - Hq - Hkv;
        __half *dst = vc + (size_t)b * kvb + (size_t)pos * Hkv * D + (size_t)hv * D;
        const float2 v = *(const float2 *)(xi + 2 * p);
        *(__half2 *)(dst + 2 * p) = __floats2half2_rn(v.x, v.y);
        return;
    }
    const int hr = rot / 2;
    int d0, d1;
    float v0, v1;
    if (p < hr) {
        if (type == 2) { d0 = 2 * p; d1 = 2 * p + 1; }
        else { d0 = p; d1 = p + hr; }
        const float a = xi[d0], bb = xi[d1];
        const float sn = sin_t[(size_t)pos * tdim + d0], cs = cos_t[(size_t)pos * tdim + d0];
        v0 = __fsub_rn(__fmul_rn(a, cs), __fmul_rn(bb, sn));
        v1 = __fadd_rn(__fmul_rn(a, sn), __fmul_rn(bb, cs));
    } else {
        d0 = rot + 2 * (p - hr); d1 = d0 + 1;
        v0 = xi[d0]; v1 = xi[d1];
    }
    if (h < Hq) {
        xi[d0] = v0; xi[d1] = v1; // each pair is owned by exactly one thread: in place is safe
    } else {
        __half *dst = kc + (size_t)b * kvb + (size_t)pos * Hkv * D + (size_t)(h - Hq) * D;
        dst[d0] = __float2half_rn(v0);
        dst[d1] = __float2half_rn(v1);
    }
}

MLLMCU_API int mllmcu_decode_rope_kv(int rope_type, float *qkv, int ld, int B, int Hq, int Hkv, int D, int rot_dim, const float *sin_t,
                                     const float *cos_t, int tdim, const int *ctl, void *kcache, void *vcache, long long kv_bstride,
                                     int cache_limit, int pdl, void *stream) {
    if ((rope_type != 2 && rope_type != 4) || D % 2 || rot_dim % 2 || rot_dim > D || rot_dim > tdim || B <= 0 || ld % 2) return -1;
    const long long total = (long long)B * (Hq + 2 * Hkv) * (D / 2);
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)((total + 255) / 256));
    cfg.blockDim = dim3(256);
    cfg.stream = (cudaStream_t)stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = pdl ? 1 : 0;
    CU_TRY(cudaLaunchKernelEx(&cfg, k_decode_rope_kv, rope_type, qkv, ld, B, Hq, Hkv, D, rot_dim, sin_t, cos_t, tdim, ctl, (__half *)kcache,
                              (__half *)vcache, kv_bstride, cache_limit));
    return 0;
}

// ------------------------------------------------------------------ decode attention
// One CTA (4 warps) per (batch row, q head, key split).  Inside a warp a group of LPK = D/16 lanes
// owns one key: each lane holds 16 of the D dims (two 16-byte fp16 loads per K row and per V row),
// the score is a 16-FMA partial + log2(LPK) shuffles, and every group runs its own online softmax
// (m, l, acc[16]) so nothing in the key loop is warp-wide.  Groups are merged by shuffles, warps
// through shared memory, splits through the workspace: the last CTA of a (row, head) to finish
// (atomic ticket) merges the nsplit partials -- no second launch.
// Arithmetic follows the reference: scores unscaled, p = expf((s - max) * (1/sqrt(D))).
struct AttnState {
    float m, l;
};
static __device__ __forceinline__ void h8_to_f(const uint4 &u, float *f) {
    const __half2 *h = (const __half2 *)&u;
#pragma unroll
    for (int i = 0; i < 4; i++) {
        const float2 t = __half22float2(h[i]);
        f[2 * i] = t.x; f[2 * i + 1] = t.y;
    }
}

// Optional fused RoPE + KV append (ROPE = true, full rotary only): q/k/v are the raw q|k|v projection row(s);
// q is rotated in registers, and the group that owns key `pos` (the token being decoded) rotates k, rounds k and v to
// fp16 exactly as the cache stores them, uses those values, and -- for the first q head of each kv head -- appends
// them to the cache.  Saves the separate rope/append launch; arithmetic identical to k_decode_rope_kv.
struct RopeArgs {
    int type, tdim, limit;
    const float *sin_t, *cos_t;
    const float *kv_src; // q|k|v rows (same row stride as q): k of kv head j at (Hq + j) * D, v at (Hq + Hkv + j) * D
    __half *kc, *vc;     // writable cache bases
};

// rotate the 16 dims [sub*16, sub*16+16) of one head held by this lane; `x` raw values.  LPK lanes hold the head.
template <int D>
static __device__ __forceinline__ void rope16(float (&x)[16], int type, int sub, const float *sn, const float *cs) {
    constexpr int LPK = D / 16;
    if (type == 2) { // pairs (2i, 2i+1): both in this lane
#pragma unroll
        for (int i = 0; i < 16; i += 2) {
            const float a = x[i], bb = x[i + 1];
            x[i] = __fsub_rn(__fmul_rn(a, cs[i]), __fmul_rn(bb, sn[i]));
            x[i + 1] = __fadd_rn(__fmul_rn(a, sn[i]), __fmul_rn(bb, cs[i]));
        }
    } else { // pairs (d, d + D/2): the partner lives LPK/2 lanes away
        const bool upper = sub >= LPK / 2;
#pragma unroll
        for (int i = 0; i < 16; i++) {
            const float other = __shfl_xor_sync(0xffffffffu, x[i], LPK / 2);
            // lower half: a*cos - b*sin (a = own, b = partner); upper half: a*sin + b*cos (a = partner, b = own)
            x[i] = upper ? __fadd_rn(__fmul_rn(other, sn[i]), __fmul_rn(x[i], cs[i])) : __fsub_rn(__fmul_rn(x[i], cs[i]), __fmul_rn(other, sn[i]));
        }
    }
}

template <int D, bool ROPE>
__global__ void __launch_bounds__(128) k_attn_decode(const float *__restrict__ q, int ldq, const __half *__restrict__ k,
                                                     const __half *__restrict__ v, float *__restrict__ o, int ldo, float *__restrict__ ws,
                                                     unsigned *__restrict__ tickets, const int *__restrict__ ctl, int Hq, int Hkv,
                                                     long long kvb, int nsplit, const RopeArgs ra) {
    constexpr int LPK = D / 16, KPW = 32 / LPK;
    __shared__ float sm[4][D + 2];
    __shared__ int s_last;
    pdl_wait();
    pdl_launch();
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int sp = blockIdx.x % nsplit, h = (blockIdx.x / nsplit) % Hq, b = blockIdx.x / (nsplit * Hq);
    const int kvh = h / (Hq / Hkv);
    const int pos = ctl[0];
    const int Sk = pos + 1;
    const int chunk = (Sk + nsplit - 1) / nsplit;
    const int c_begin = sp * chunk, c_end = min(Sk, c_begin + chunk);
    const int g = lane / LPK, sub = lane % LPK;
    const float scale = __fdiv_rn(1.0f, __fsqrt_rn((float)D));

    float qv[16], acc[16];
    {
        const float4 *qp = (const float4 *)(q + (size_t)b * ldq + (size_t)h * D + sub * 16);
#pragma unroll
        for (int i = 0; i < 4; i++) { const float4 t = qp[i]; qv[4 * i] = t.x; qv[4 * i + 1] = t.y; qv[4 * i + 2] = t.z; qv[4 * i + 3] = t.w; }
    }
    // table row of this position for the lane's 16 dims (loaded where used: keeping 32 values live across the key loop costs occupancy)
    auto load_sincos = [&](float (&sn)[16], float (&cs)[16]) {
        const float4 *sp4 = (const float4 *)(ra.sin_t + (size_t)pos * ra.tdim + sub * 16), *cp4 = (const float4 *)(ra.cos_t + (size_t)pos * ra.tdim + sub * 16);
#pragma unroll
        for (int i = 0; i < 4; i++) {
            const float4 a = sp4[i], c = cp4[i];
            sn[4 * i] = a.x; sn[4 * i + 1] = a.y; sn[4 * i + 2] = a.z; sn[4 * i + 3] = a.w;
            cs[4 * i] = c.x; cs[4 * i + 1] = c.y; cs[4 * i + 2] = c.z; cs[4 * i + 3] = c.w;
        }
    };
    if (ROPE) {
        if (pos < 0 || pos >= ra.limit) return; // the host refuses such a step; never touch memory outside the cache
        float sn[16], cs[16];
        load_sincos(sn, cs);
        rope16<D>(qv, ra.type, sub, sn, cs);
    }
#pragma unroll
    for (int i = 0; i < 16; i++) acc[i] = 0.f;
    float m = -INFINITY, l = 0.f;
    const size_t rs = (size_t)Hkv * D;
    const __half *kb = k + (size_t)b * kvb + (size_t)kvh * D + sub * 16;
    const __half *vb = v + (size_t)b * kvb + (size_t)kvh * D + sub * 16;
#pragma unroll 2
    for (int c0 = c_begin + warp * KPW; c0 < c_end; c0 += 4 * KPW) { // warp-uniform trip count
        const int c = c0 + g;
        const bool ok = c < c_end;
        const bool fresh = ROPE && ((c0 + KPW > pos) && (c0 <= pos)); // warp-uniform: this pass contains the new token's key
        const size_t off = (size_t)((ok && !(ROPE && c == pos)) ? c : c_begin) * rs;
        uint4 k0 = *(const uint4 *)(kb + off), k1 = *(const uint4 *)(kb + off + 8);
        uint4 v0 = *(const uint4 *)(vb + off), v1 = *(const uint4 *)(vb + off + 8);
        if (fresh) { // every lane takes part in the shuffles of rope16; only the owning group keeps the result
            float kx[16], vx[16];
            const float4 *kp = (const float4 *)(ra.kv_src + (size_t)b * ldq + (size_t)(Hq + kvh) * D + sub * 16);
            const float4 *vp = (const float4 *)(ra.kv_src + (size_t)b * ldq + (size_t)(Hq + Hkv + kvh) * D + sub * 16);
#pragma unroll
            for (int i = 0; i < 4; i++) {
                const float4 t = kp[i], u = vp[i];
                kx[4 * i] = t.x; kx[4 * i + 1] = t.y; kx[4 * i + 2] = t.z; kx[4 * i + 3] = t.w;
                vx[4 * i] = u.x; vx[4 * i + 1] = u.y; vx[4 * i + 2] = u.z; vx[4 * i + 3] = u.w;
            }
            {
                float sn[16], cs[16];
                load_sincos(sn, cs);
                rope16<D>(kx, ra.type, sub, sn, cs);
            }
            if (ok && c == pos) {
                __half2 *kh = (__half2 *)&k0, *kh1 = (__half2 *)&k1, *vh = (__half2 *)&v0, *vh1 = (__half2 *)&v1;
#pragma unroll
                for (int i = 0; i < 4; i++) {
                    kh[i] = __floats2half2_rn(kx[2 * i], kx[2 * i + 1]); kh1[i] = __floats2half2_rn(kx[8 + 2 * i], kx[8 + 2 * i + 1]);
                    vh[i] = __floats2half2_rn(vx[2 * i], vx[2 * i + 1]); vh1[i] = __floats2half2_rn(vx[8 + 2 * i], vx[8 + 2 * i + 1]);
                }
                if (h % (Hq / Hkv) == 0) { // one writer per kv head
                    __half *kd = ra.kc + (size_t)b * kvb + (size_t)pos * rs + (size_t)kvh * D + sub * 16;
                    __half *vd = ra.vc + (size_t)b * kvb + (size_t)pos * rs + (size_t)kvh * D + sub * 16;
                    *(uint4 *)kd = k0; *(uint4 *)(kd + 8) = k1;
                    *(uint4 *)vd = v0; *(uint4 *)(vd + 8) = v1;
                }
            }
        }
        float kf[16], vf[16];
        h8_to_f(k0, kf); h8_to_f(k1, kf + 8);
        float s = 0.f;
#pragma unroll
        for (int i = 0; i < 16; i++) s = fmaf(qv[i], kf[i], s);
#pragma unroll
        for (int x = 1; x < LPK; x <<= 1) s += __shfl_xor_sync(0xffffffffu, s, x);
        if (ok) {
            h8_to_f(v0, vf); h8_to_f(v1, vf + 8);
            const float m_new = fmaxf(m, s);
            const float sc = expf((m - m_new) * scale);
            const float p = expf((s - m_new) * scale);
            l = fmaf(l, sc, p);
#pragma unroll
            for (int i = 0; i < 16; i++) acc[i] = fmaf(p, vf[i], acc[i] * sc);
            m = m_new;
        }
    }
    // ---- merge the KPW key groups of the warp (lanes with equal `sub`)
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
    __syncthreads();
    // ---- merge the 4 warps: thread t < D owns output dim t
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
    float *op = o + (size_t)b * ldo + (size_t)h * D;
    if (nsplit == 1) {
        if (tid < D) op[tid] = at * (1.0f / lt);
        return;
    }
    const size_t row = (size_t)b * Hq + h;
    float *wp = ws + (row * nsplit + sp) * (D + 2);
    if (tid < D) {
        if (tid == 0) { wp[0] = mt; wp[1] = lt; }
        wp[2 + tid] = at;
    }
    __threadfence();
    __syncthreads();
    if (tid == 0) {
        const unsigned t = atomicAdd(&tickets[row], 1u);
        s_last = (t == (unsigned)nsplit - 1u);
        if (s_last) tickets[row] = 0; // ready for the next launch
    }
    __syncthreads();
    if (!s_last) return;
    __threadfence();
    if (tid < D) {
        const float *wr = ws + row * nsplit * (D + 2);
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

MLLMCU_API int mllmcu_decode_attn_nsplit(int B, int Hq) {
    // ~2 CTAs of 4 warps per SM (swept on B200: 1x 1.70, 2x 1.66, 4x 1.82 ms per 7B decode step)
    static int mult = 0;
    if (mult == 0) { const char *e = getenv("MLLMCU_ATTN_SPLIT_MULT"); mult = (e && e[0]) ? atoi(e) : 2; if (mult < 1) mult = 1; }
    int n = (mult * 148) / (B * Hq);
    if (n < 1) n = 1;
    if (n > 32) n = 32;
    return n;
}
MLLMCU_API size_t mllmcu_decode_attn_workspace_bytes(int B, int Hq, int D) {
    const size_t rows = (size_t)B * Hq;
    return rows * mllmcu_decode_attn_nsplit(B, Hq) * (D + 2) * sizeof(float) + rows * sizeof(unsigned) + 256;
}

template <int D, bool ROPE>
static int launch_attn_decode(const float *q, int ldq, const __half *k, const __half *v, float *o, int ldo, float *ws, unsigned *tickets,
                              const int *ctl, int B, int Hq, int Hkv, long long kvb, int nsplit, const RopeArgs &ra, int pdl, cudaStream_t st) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)(B * Hq * nsplit));
    cfg.blockDim = dim3(128);
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = pdl ? 1 : 0;
    CU_TRY(cudaLaunchKernelEx(&cfg, k_attn_decode<D, ROPE>, q, ldq, k, v, o, ldo, ws, tickets, ctl, Hq, Hkv, kvb, nsplit, ra));
    return 0;
}

static int decode_attn_common(const float *q, int ldq, const void *k, const void *v, float *o, int ldo, const int *ctl, int B, int Hq, int Hkv,
                              int D, long long kv_bstride, void *ws, size_t ws_bytes, const RopeArgs *ra, int pdl, void *stream) {
    if (B <= 0 || Hq <= 0 || Hkv <= 0 || Hq % Hkv || (ldq & 3) || ((uintptr_t)q & 15) || ((uintptr_t)k & 15) || ((uintptr_t)v & 15)) return -1;
    if (D != 32 && D != 64 && D != 128) return -2;
    if ((kv_bstride & 7) || ((Hkv * D) & 7)) return -2;
    if (ws == nullptr || ws_bytes < mllmcu_decode_attn_workspace_bytes(B, Hq, D)) return -3;
    const int nsplit = mllmcu_decode_attn_nsplit(B, Hq);
    float *wsf = (float *)ws;
    unsigned *tickets = (unsigned *)(wsf + (size_t)B * Hq * nsplit * (D + 2));
    cudaStream_t st = (cudaStream_t)stream;
    const __half *kh = (const __half *)k, *vh = (const __half *)v;
    RopeArgs none = {};
    if (ra) {
        if (D == 128) return launch_attn_decode<128, true>(q, ldq, kh, vh, o, ldo, wsf, tickets, ctl, B, Hq, Hkv, kv_bstride, nsplit, *ra, pdl, st);
        if (D == 64) return launch_attn_decode<64, true>(q, ldq, kh, vh, o, ldo, wsf, tickets, ctl, B, Hq, Hkv, kv_bstride, nsplit, *ra, pdl, st);
        return launch_attn_decode<32, true>(q, ldq, kh, vh, o, ldo, wsf, tickets, ctl, B, Hq, Hkv, kv_bstride, nsplit, *ra, pdl, st);
    }
    if (D == 128) return launch_attn_decode<128, false>(q, ldq, kh, vh, o, ldo, wsf, tickets, ctl, B, Hq, Hkv, kv_bstride, nsplit, none, pdl, st);
    if (D == 64) return launch_attn_decode<64, false>(q, ldq, kh, vh, o, ldo, wsf, tickets, ctl, B, Hq, Hkv, kv_bstride, nsplit, none, pdl, st);
    return launch_attn_decode<32, false>(q, ldq, kh, vh, o, ldo, wsf, tickets, ctl, B, Hq, Hkv, kv_bstride, nsplit, none, pdl, st);
}

// q [B][ldq] fp32 (head h at h*D), o [B][ldo]; K/V fp16 at b*kv_bstride + s*Hkv*D + h*D; the KV length is ctl[0]+1.
// ws: mllmcu_decode_attn_workspace_bytes, zero-initialised once (the ticket counters live at its end and reset themselves).
MLLMCU_API int mllmcu_decode_attn(const float *q, int ldq, const void *k, const void *v, float *o, int ldo, const int *ctl, int B, int Hq,
                                  int Hkv, int D, long long kv_bstride, void *ws, size_t ws_bytes, int pdl, void *stream) {
    return decode_attn_common(q, ldq, k, v, o, ldo, ctl, B, Hq, Hkv, D, kv_bstride, ws, ws_bytes, nullptr, pdl, stream);
}

// Same with q_rope / k_rope / k_cache / v_cache fused in: qkv [B][ld] holds the RAW q|k|v projection rows; the kernel rotates q
// in registers, rotates k, appends k and v (fp16) at cache row ctl[0] and attends over ctl[0]+1 keys.  Full rotary only
// (rot_dim == D == tdim); qkv is not modified.
MLLMCU_API int mllmcu_decode_attn_rope(int rope_type, const float *qkv, int ld, void *kcache, void *vcache, float *o, int ldo, const int *ctl,
                                       int B, int Hq, int Hkv, int D, const float *sin_t, const float *cos_t, int tdim, long long kv_bstride,
                                       int cache_limit, void *ws, size_t ws_bytes, int pdl, void *stream) {
    if ((rope_type != 2 && rope_type != 4) || tdim != D || ((uintptr_t)sin_t & 15) || ((uintptr_t)cos_t & 15)) return -2;
    RopeArgs ra = {rope_type, tdim, cache_limit, sin_t, cos_t, qkv, (__half *)kcache, (__half *)vcache};
    return decode_attn_common(qkv, ld, kcache, vcache, o, ldo, ctl, B, Hq, Hkv, D, kv_bstride, ws, ws_bytes, &ra, pdl, stream);
}

// ------------------------------------------------------------------ greedy pick with token log
// ids_out[b] (fp32, the next step's input_ids) and tok_log[ctl[0] * rows + b] (int32): the token chosen
// after consuming position ctl[0].  First maximum wins (Generate.cpp:46-53).
__global__ void __launch_bounds__(1024) k_decode_argmax(const float *__restrict__ x, int n, float *__restrict__ ids_out, int *__restrict__ tok_log,
                                                        const int *__restrict__ ctl, int log_rows) {
    const float *xr = x + (size_t)blockIdx.x * n;
    float best = -INFINITY;
    int bi = 0x7fffffff;
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
        const float v = xr[i];
        if (v > best || (v == best && i < bi)) { best = v; bi = i; }
    }
    __shared__ float sv[32];
    __shared__ int si[32];
    for (int round = 0; round < 2; round++) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            const float ov = __shfl_xor_sync(0xffffffffu, best, o);
            const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
            if (ov > best || (ov == best && oi < bi)) { best = ov; bi = oi; }
        }
        if (round == 1) break;
        if ((threadIdx.x & 31) == 0) { sv[threadIdx.x >> 5] = best; si[threadIdx.x >> 5] = bi; }
        __syncthreads();
        if (threadIdx.x >= 32) return;
        const int nw = blockDim.x >> 5;
        best = threadIdx.x < nw ? sv[threadIdx.x] : -INFINITY;
        bi = threadIdx.x < nw ? si[threadIdx.x] : 0x7fffffff;
    }
    if (threadIdx.x == 0) {
        const int r = (bi == 0x7fffffff) ? 0 : bi;
        ids_out[blockIdx.x] = (float)r;
        if (tok_log) {
            const int pos = ctl[0];
            if (pos >= 0 && pos < log_rows) tok_log[(size_t)pos * gridDim.x + blockIdx.x] = r;
        }
    }
}
MLLMCU_API int mllmcu_decode_argmax(const float *logits, int rows, int n, float *ids_out, int *tok_log, const int *ctl, int log_rows,
                                    void *stream) {
    if (rows <= 0 || n <= 0 || ids_out == nullptr || (tok_log && ctl == nullptr)) return -1;
    k_decode_argmax<<<rows, 1024, 0, (cudaStream_t)stream>>>(logits, n, ids_out, tok_log, ctl, log_rows);
    CU_LAUNCH_CHECK();
    return 0;
}
