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
        v0 = rope_lo(a, bb, sn, cs);
        v1 = rope_hi(a, bb, sn, cs);
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
// One CTA (ATTN_WARPS warps) per (batch row, q head, key split).  Inside a warp a group of LPK = D/16 lanes
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
    unsigned long long *prof; // development aid (MLLMCU_ATTN_PROF=1): 8 %globaltimer slots per launch
    int early;                // ctl[0] is final before the dependency resolves: position, K/V rows and the table row are read ahead of the wait
};

// ---- thread-block cluster helpers: the key splits of one (row, head) form a cluster and merge through rank 0's shared memory
static __device__ __forceinline__ uint32_t smem_addr_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
static __device__ __forceinline__ uint32_t cluster_map_rank(uint32_t saddr, uint32_t rank) {
    uint32_t r;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(saddr), "r"(rank));
    return r;
}
static __device__ __forceinline__ void cluster_st_f32(uint32_t addr, float v) { asm volatile("st.shared::cluster.f32 [%0], %1;" ::"r"(addr), "f"(v) : "memory"); }
static __device__ __forceinline__ void cluster_sync_all() {
    asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
    asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}

// rotate the 16 dims [sub*16, sub*16+16) of one head held by this lane; `x` raw values.  LPK lanes hold the head.
template <int D>
static __device__ __forceinline__ void rope16(float (&x)[16], int type, int sub, const float *sn, const float *cs) {
    constexpr int LPK = D / 16;
    if (type == 2) { // pairs (2i, 2i+1): both in this lane
#pragma unroll
        for (int i = 0; i < 16; i += 2) {
            const float a = x[i], bb = x[i + 1];
            x[i] = rope_lo(a, bb, sn[i], cs[i]);
            x[i + 1] = rope_hi(a, bb, sn[i], cs[i]);
        }
    } else { // pairs (d, d + D/2): the partner lives LPK/2 lanes away
        const bool upper = sub >= LPK / 2;
#pragma unroll
        for (int i = 0; i < 16; i++) {
            const float other = __shfl_xor_sync(0xffffffffu, x[i], LPK / 2);
            // lower half: a*cos - b*sin (a = own, b = partner); upper half: a*sin + b*cos (a = partner, b = own)
            x[i] = upper ? rope_hi(other, x[i], sn[i], cs[i]) : rope_lo(x[i], other, sn[i], cs[i]);
        }
    }
}

#define ATTN_PROF_RING 4096
static unsigned long long *g_attn_prof = nullptr;
static int g_attn_prof_on = -1, g_attn_prof_n = 0;
#define ATTN_WARPS 4 // warps per CTA: 128 threads x <= 200 registers sit beside a resident Linear CTA (512 x 64) on the same SM
#define ATTN_NB 6    // key passes whose K/V loads are in flight together (registers: 4 x uint4 per pass): 96 keys per CTA before any refill
#define ATTN_MAXS 16 // splits merged from registers in one L2 round trip (more: a second pass over the workspace)

template <int D, bool ROPE, bool CLUSTER>
__global__ void __launch_bounds__(ATTN_WARPS * 32) k_attn_decode(const float *__restrict__ q, int ldq, const __half *__restrict__ k,
                                                                const __half *__restrict__ v, float *__restrict__ o, int ldo, float *__restrict__ ws,
                                                                unsigned *__restrict__ tickets, const int *__restrict__ ctl, int Hq, int Hkv,
                                                                long long kvb, int nsplit, const RopeArgs ra) {
    constexpr int LPK = D / 16, KPW = 32 / LPK, NW = ATTN_WARPS, KPP = NW * KPW, NB = ATTN_NB;
    __shared__ float sm[NW][D + 2];
    __shared__ int s_last;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int sp = blockIdx.x % nsplit, h = (blockIdx.x / nsplit) % Hq, b = blockIdx.x / (nsplit * Hq);
    const int kvh = h / (Hq / Hkv);
    const size_t rs = (size_t)Hkv * D;
#define ATTN_STAMP(k) do { if (ROPE && ra.prof && blockIdx.x == 0 && tid == 0) { unsigned long long t_; asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t_)); ra.prof[k] = t_; } } while (0)
    ATTN_STAMP(0);
    pdl_launch();
    const int g = lane / LPK, sub = lane % LPK;
    const float scale = __fdiv_rn(1.0f, __fsqrt_rn((float)D));
    const __half *kb = k + (size_t)b * kvb + (size_t)kvh * D + sub * 16;
    const __half *vb = v + (size_t)b * kvb + (size_t)kvh * D + sub * 16;
    const bool early = ROPE && ra.early; // kernel-uniform
    // [c_begin, c_end): this split's keys; [c_begin, e_end): those already in the cache (with ROPE the last key, `pos`, is the
    // token being decoded: it is rotated and appended by warp 0 of the split that owns it, apart from the cached keys)
    int pos = 0, c_begin = 0, c_end = 0, e_end = 0;
    auto take_pos = [&]() {
        pos = ctl[0];
        const int Sk = pos + 1, chunk = (Sk + nsplit - 1) / nsplit;
        c_begin = sp * chunk;
        c_end = min(Sk, c_begin + chunk);
        e_end = ROPE ? min(c_end, pos) : c_end;
    };
    // K/V rows of NB passes are in flight together; lanes past the end read the split's first row (finite values, masked later)
    uint4 kk[NB][2], vv[NB][2];
#pragma unroll
    for (int j = 0; j < NB; j++) kk[j][0] = kk[j][1] = vv[j][0] = vv[j][1] = make_uint4(0u, 0u, 0u, 0u);
    auto load_k = [&](int j, int c0) {
        const int c = c0 + g;
        const size_t off = (size_t)(c < e_end ? c : c_begin) * rs;
        kk[j][0] = *(const uint4 *)(kb + off); kk[j][1] = *(const uint4 *)(kb + off + 8);
    };
    auto load_v = [&](int j, int c0) {
        const int c = c0 + g;
        const size_t off = (size_t)(c < e_end ? c : c_begin) * rs;
        vv[j][0] = *(const uint4 *)(vb + off); vv[j][1] = *(const uint4 *)(vb + off + 8);
    };
    auto issue_loads = [&](int cb0) {
#pragma unroll
        for (int j = 0; j < NB; j++) {
            const int c0 = cb0 + j * KPP;
            if (c0 >= e_end) break; // warp-uniform
            load_k(j, c0);
            load_v(j, c0);
        }
    };
    // table row of this position for the lane's 16 dims (shared by q and the new key)
    auto load_sincos = [&](float (&sn)[16], float (&cs)[16]) {
        const float4 *sp4 = (const float4 *)(ra.sin_t + (size_t)pos * ra.tdim + sub * 16), *cp4 = (const float4 *)(ra.cos_t + (size_t)pos * ra.tdim + sub * 16);
#pragma unroll
        for (int i = 0; i < 4; i++) {
            const float4 a = sp4[i], c = cp4[i];
            sn[4 * i] = a.x; sn[4 * i + 1] = a.y; sn[4 * i + 2] = a.z; sn[4 * i + 3] = a.w;
            cs[4 * i] = c.x; cs[4 * i + 1] = c.y; cs[4 * i + 2] = c.z; cs[4 * i + 3] = c.w;
        }
    };
    float qsn[16], qcs[16];
    bool bad_pos = false;
    if (early) {
        // The caller guarantees ctl[0] is final (an earlier kernel of the chain triggered its dependents only after ITS wait), and
        // the cache rows of earlier tokens were complete before this step began: position, K/V rows and the table row are all
        // read while the q|k|v projection in front is still running.
        take_pos();
        bad_pos = pos < 0 || pos >= ra.limit;
        if (!bad_pos) {
            issue_loads(c_begin + warp * KPW);
            load_sincos(qsn, qcs);
            for (int c = c_begin + NB * KPP + tid; c < e_end; c += NW * 32) { // keys beyond the register window: into L2
                const __half *kr = k + (size_t)b * kvb + (size_t)c * rs + (size_t)kvh * D, *vr = v + (size_t)b * kvb + (size_t)c * rs + (size_t)kvh * D;
                asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(kr), "r"(D * 2) : "memory");
                asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(vr), "r"(D * 2) : "memory");
            }
        }
    } else if (ROPE) {
        // ctl[0] may still be the previous token's position while the embedding kernel of this step is in flight: only ask L2
        // for this CTA's slice of the cache (a slice that is off by one key prefetches a neighbour row, nothing more)
        const int p0 = min(max(ctl[0], 0), ra.limit - 1);
        const int ch0 = (p0 + nsplit) / nsplit, cb = sp * ch0, ce = min(p0, cb + ch0);
        for (int c = cb + tid; c < ce; c += NW * 32) {
            const __half *kr = k + (size_t)b * kvb + (size_t)c * rs + (size_t)kvh * D, *vr = v + (size_t)b * kvb + (size_t)c * rs + (size_t)kvh * D;
            asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(kr), "r"(D * 2) : "memory");
            asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(vr), "r"(D * 2) : "memory");
        }
    }
    pdl_wait();
    ATTN_STAMP(1);
    if (!early) {
        take_pos();
        bad_pos = ROPE && (pos < 0 || pos >= ra.limit);
    }
    if (bad_pos) { // the host refuses such a step; never touch memory outside the cache (uniform over the grid)
        if (CLUSTER) cluster_sync_all();
        return;
    }
    const int first = c_begin + warp * KPW;
    if (!early) {
        issue_loads(first);
        if (ROPE) load_sincos(qsn, qcs);
    }

    float qv[16], acc[16];
    {
        const float4 *qp = (const float4 *)(q + (size_t)b * ldq + (size_t)h * D + sub * 16);
#pragma unroll
        for (int i = 0; i < 4; i++) { const float4 t = qp[i]; qv[4 * i] = t.x; qv[4 * i + 1] = t.y; qv[4 * i + 2] = t.z; qv[4 * i + 3] = t.w; }
    }
    // the raw k and v of the new token, fetched together with q by the one warp that will need them
    const bool fresh = ROPE && warp == 0 && pos >= c_begin && pos < c_end; // warp-uniform
    float kx[16], vx[16];
    if (fresh) {
        const float4 *kp = (const float4 *)(ra.kv_src + (size_t)b * ldq + (size_t)(Hq + kvh) * D + sub * 16);
        const float4 *vp = (const float4 *)(ra.kv_src + (size_t)b * ldq + (size_t)(Hq + Hkv + kvh) * D + sub * 16);
#pragma unroll
        for (int i = 0; i < 4; i++) {
            const float4 t = kp[i], u = vp[i];
            kx[4 * i] = t.x; kx[4 * i + 1] = t.y; kx[4 * i + 2] = t.z; kx[4 * i + 3] = t.w;
            vx[4 * i] = u.x; vx[4 * i + 1] = u.y; vx[4 * i + 2] = u.z; vx[4 * i + 3] = u.w;
        }
    }
    if (ROPE) rope16<D>(qv, ra.type, sub, qsn, qcs);
#pragma unroll
    for (int i = 0; i < 16; i++) acc[i] = 0.f;
    float m = -INFINITY, l = 0.f;
    if (ROPE && ra.prof && blockIdx.x == 0 && tid == 0) { unsigned long long t_; asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t_)); ra.prof[2] = t_ + (qv[0] == 1.2345e-30f ? 1 : 0); }
    if (fresh) {
        // every group of the warp rotates the key (the shuffles of rope16 need whole groups); group 0 keeps it as its first key.
        // k and v are rounded to fp16 exactly as the cache stores them, and appended by the first q head of each kv head.
        rope16<D>(kx, ra.type, sub, qsn, qcs);
        uint4 k0, k1, v0, v1;
        __half2 *kh = (__half2 *)&k0, *kh1 = (__half2 *)&k1, *vh = (__half2 *)&v0, *vh1 = (__half2 *)&v1;
#pragma unroll
        for (int i = 0; i < 4; i++) {
            kh[i] = __floats2half2_rn(kx[2 * i], kx[2 * i + 1]); kh1[i] = __floats2half2_rn(kx[8 + 2 * i], kx[8 + 2 * i + 1]);
            vh[i] = __floats2half2_rn(vx[2 * i], vx[2 * i + 1]); vh1[i] = __floats2half2_rn(vx[8 + 2 * i], vx[8 + 2 * i + 1]);
        }
        if (g == 0 && h % (Hq / Hkv) == 0) { // one writer per kv head
            __half *kd = ra.kc + (size_t)b * kvb + (size_t)pos * rs + (size_t)kvh * D + sub * 16;
            __half *vd = ra.vc + (size_t)b * kvb + (size_t)pos * rs + (size_t)kvh * D + sub * 16;
            *(uint4 *)kd = k0; *(uint4 *)(kd + 8) = k1;
            *(uint4 *)vd = v0; *(uint4 *)(vd + 8) = v1;
        }
        float kf[16];
        h8_to_f(k0, kf); h8_to_f(k1, kf + 8);
        float s = 0.f;
#pragma unroll
        for (int i = 0; i < 16; i++) s = fmaf(qv[i], kf[i], s);
#pragma unroll
        for (int x = 1; x < LPK; x <<= 1) s += __shfl_xor_sync(0xffffffffu, s, x);
        if (g == 0) { // first key of this group: weight exp(0) = 1
            m = s; l = 1.0f;
            h8_to_f(v0, acc); h8_to_f(v1, acc + 8);
        }
    }
    for (int cb0 = first; cb0 < e_end; cb0 += NB * KPP) { // warp-uniform trip count
        // scores of the NB passes first (independent chains), then ONE softmax rescale for the batch
        float sj[NB];
#pragma unroll
        for (int j = 0; j < NB; j++) {
            float kf[16];
            h8_to_f(kk[j][0], kf); h8_to_f(kk[j][1], kf + 8);
            float s0 = 0.f, s1 = 0.f, s2 = 0.f, s3 = 0.f;
#pragma unroll
            for (int i = 0; i < 4; i++) {
                s0 = fmaf(qv[i], kf[i], s0); s1 = fmaf(qv[4 + i], kf[4 + i], s1);
                s2 = fmaf(qv[8 + i], kf[8 + i], s2); s3 = fmaf(qv[12 + i], kf[12 + i], s3);
            }
            sj[j] = (s0 + s1) + (s2 + s3);
            const int cn0 = cb0 + (j + NB) * KPP;
            if (cn0 < e_end) load_k(j, cn0); // this slot's K is consumed: refill it with the pass NB ahead (warp-uniform)
        }
#pragma unroll
        for (int x = 1; x < LPK; x <<= 1)
#pragma unroll
            for (int j = 0; j < NB; j++) sj[j] += __shfl_xor_sync(0xffffffffu, sj[j], x);
        float bm = -INFINITY;
#pragma unroll
        for (int j = 0; j < NB; j++) {
            if (cb0 + j * KPP + g >= e_end) sj[j] = -INFINITY;
            bm = fmaxf(bm, sj[j]);
        }
        const float m_new = fmaxf(m, bm);
        const float sc = (m == -INFINITY) ? 0.f : expf((m - m_new) * scale);
        float pj[NB], psum = 0.f;
#pragma unroll
        for (int j = 0; j < NB; j++) {
            pj[j] = (sj[j] == -INFINITY) ? 0.f : expf((sj[j] - m_new) * scale);
            psum += pj[j];
        }
        l = fmaf(l, sc, psum);
#pragma unroll
        for (int i = 0; i < 16; i++) acc[i] *= sc;
#pragma unroll
        for (int j = 0; j < NB; j++) {
            float vf[16];
            h8_to_f(vv[j][0], vf); h8_to_f(vv[j][1], vf + 8);
#pragma unroll
            for (int i = 0; i < 16; i++) acc[i] = fmaf(pj[j], vf[i], acc[i]);
            const int cn0 = cb0 + (j + NB) * KPP;
            if (cn0 < e_end) load_v(j, cn0);
        }
        m = m_new;
    }
    if (ROPE && ra.prof && blockIdx.x == 0 && tid == 0) { unsigned long long t_; asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t_)); ra.prof[3] = t_ + (acc[0] == 1.2345e-30f ? 1 : 0); }
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
    // ---- merge the warps: thread t < D owns output dim t
    float mt = -INFINITY, lt = 0.f, at = 0.f;
    if (tid < D) {
#pragma unroll
        for (int w = 0; w < NW; w++) mt = fmaxf(mt, sm[w][0]);
#pragma unroll
        for (int w = 0; w < NW; w++) {
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
    if (CLUSTER) {
        // the nsplit CTAs of this (row, head) are one cluster (rank == sp): partials go straight into rank 0's shared memory,
        // one cluster barrier orders them, rank 0 merges -- no global round trip, no ticket
        __shared__ float part[8][D + 2];
        const uint32_t dst = cluster_map_rank(smem_addr_u32(&part[sp][0]), 0);
        if (tid < D) {
            if (tid == 0) { cluster_st_f32(dst, mt); cluster_st_f32(dst + 4, lt); }
            cluster_st_f32(dst + 8 + tid * 4, at);
        }
        ATTN_STAMP(4);
        cluster_sync_all();
        ATTN_STAMP(5);
        if (sp != 0) return;
        if (ROPE && ra.prof && row == 0 && tid == 0) { unsigned long long t_; asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t_)); ra.prof[6] = t_; }
        if (tid < D) {
            float mm = -INFINITY, ll = 0.f, aa = 0.f;
            for (int s2 = 0; s2 < nsplit; s2++) mm = fmaxf(mm, part[s2][0]);
            for (int s2 = 0; s2 < nsplit; s2++) {
                const float ms = part[s2][0];
                const float f = (ms == -INFINITY) ? 0.f : expf((ms - mm) * scale);
                ll = fmaf(part[s2][1], f, ll);
                aa = fmaf(part[s2][2 + tid], f, aa);
            }
            op[tid] = aa * (1.0f / ll);
        }
        if (ROPE && ra.prof && row == 0 && tid == 0) { unsigned long long t_; asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t_)); ra.prof[7] = t_; }
        return;
    }
    float *wp = ws + (row * nsplit + sp) * (D + 2);
    if (tid < D) {
        if (tid == 0) { wp[0] = mt; wp[1] = lt; }
        wp[2 + tid] = at;
    }
    __syncthreads();
    ATTN_STAMP(4);
    if (tid == 0) {
        // release the CTA's partial (ordered before this point by the barrier) and acquire the others' in one atomic
        unsigned t;
        asm volatile("atom.acq_rel.gpu.global.add.u32 %0, [%1], 1;" : "=r"(t) : "l"(tickets + row) : "memory");
        s_last = (t == (unsigned)nsplit - 1u);
        if (s_last) tickets[row] = 0; // ready for the next launch
    }
    __syncthreads();
    ATTN_STAMP(5);
    if (!s_last) return;
    if (ROPE && ra.prof && row == 0 && tid == 0) { unsigned long long t_; asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t_)); ra.prof[6] = t_; }
    if (tid < D) {
        const float *wr = ws + row * nsplit * (D + 2);
        float mm = -INFINITY, ll = 0.f, aa = 0.f;
        if (nsplit <= ATTN_MAXS) { // every partial in flight at once
            float ms[ATTN_MAXS], ls[ATTN_MAXS], as[ATTN_MAXS];
#pragma unroll
            for (int s = 0; s < ATTN_MAXS; s++) {
                const bool on = s < nsplit;
                const float *r = wr + (size_t)(on ? s : 0) * (D + 2);
                ms[s] = on ? __ldcg(r) : -INFINITY;
                ls[s] = __ldcg(r + 1);
                as[s] = __ldcg(r + 2 + tid);
            }
#pragma unroll
            for (int s = 0; s < ATTN_MAXS; s++) mm = fmaxf(mm, ms[s]);
#pragma unroll
            for (int s = 0; s < ATTN_MAXS; s++) {
                const float f = (ms[s] == -INFINITY) ? 0.f : expf((ms[s] - mm) * scale);
                ll = fmaf(ls[s], f, ll);
                aa = fmaf(as[s], f, aa);
            }
        } else {
            for (int s = 0; s < nsplit; s++) mm = fmaxf(mm, __ldcg(wr + (size_t)s * (D + 2)));
            for (int s = 0; s < nsplit; s++) {
                const float ms = __ldcg(wr + (size_t)s * (D + 2));
                if (ms == -INFINITY) continue;
                const float f = expf((ms - mm) * scale);
                ll = fmaf(__ldcg(wr + (size_t)s * (D + 2) + 1), f, ll);
                aa = fmaf(__ldcg(wr + (size_t)s * (D + 2) + 2 + tid), f, aa);
            }
        }
        op[tid] = aa * (1.0f / ll);
    }
    if (ROPE && ra.prof && row == 0 && tid == 0) { unsigned long long t_; asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t_)); ra.prof[7] = t_; }
}

static bool attn_use_cluster() {
    static int on = -1;
    if (on < 0) { const char *e = getenv("MLLMCU_ATTN_NO_CLUSTER"); on = (e && e[0] && e[0] != '0') ? 0 : 1; }
    return on == 1;
}
MLLMCU_API int mllmcu_decode_attn_nsplit(int B, int Hq) {
    // ~2 CTAs of 4 warps per SM (swept on B200: 1x 1.70, 2x 1.66, 4x 1.82 ms per 7B decode step)
    static int mult = 0;
    if (mult == 0) { const char *e = getenv("MLLMCU_ATTN_SPLIT_MULT"); mult = (e && e[0]) ? atoi(e) : 2; if (mult < 1) mult = 1; }
    int n = (mult * cu_sm_count()) / (B * Hq);
    if (n < 1) n = 1;
    if (n > 32) n = 32;
    if (attn_use_cluster() && n > 8) n = 8; // the splits of one head form a thread-block cluster (portable maximum 8)
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
    cfg.blockDim = dim3(ATTN_WARPS * 32);
    cfg.stream = st;
    cudaLaunchAttribute attr[2];
    int na = 0;
    if (pdl) {
        attr[na].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        attr[na].val.programmaticStreamSerializationAllowed = 1;
        na++;
    }
    const bool cluster = attn_use_cluster() && nsplit > 1 && nsplit <= 8;
    if (cluster) {
        attr[na].id = cudaLaunchAttributeClusterDimension;
        attr[na].val.clusterDim.x = (unsigned)nsplit;
        attr[na].val.clusterDim.y = 1;
        attr[na].val.clusterDim.z = 1;
        na++;
    }
    cfg.attrs = attr;
    cfg.numAttrs = na;
    if (cluster) CU_TRY(cudaLaunchKernelEx(&cfg, k_attn_decode<D, ROPE, true>, q, ldq, k, v, o, ldo, ws, tickets, ctl, Hq, Hkv, kvb, nsplit, ra));
    else CU_TRY(cudaLaunchKernelEx(&cfg, k_attn_decode<D, ROPE, false>, q, ldq, k, v, o, ldo, ws, tickets, ctl, Hq, Hkv, kvb, nsplit, ra));
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
    RopeArgs ra = {rope_type, tdim, cache_limit, sin_t, cos_t, qkv, (__half *)kcache, (__half *)vcache, nullptr, pdl == 2 ? 1 : 0};
    if (g_attn_prof_on < 0) {
        const char *e = getenv("MLLMCU_ATTN_PROF");
        g_attn_prof_on = (e && e[0] == '1') ? 1 : 0;
        if (g_attn_prof_on && (cudaMalloc(&g_attn_prof, ATTN_PROF_RING * 8 * 8) != cudaSuccess || cudaMemset(g_attn_prof, 0, ATTN_PROF_RING * 8 * 8) != cudaSuccess)) g_attn_prof_on = 0;
    }
    if (g_attn_prof_on) { ra.prof = g_attn_prof + (size_t)(g_attn_prof_n % ATTN_PROF_RING) * 8; g_attn_prof_n++; }
    return decode_attn_common(qkv, ld, kcache, vcache, o, ldo, ctl, B, Hq, Hkv, D, kv_bstride, ws, ws_bytes, &ra, pdl, stream);
}

// development aid: the last n fused-rope attention launches' stamps, oldest first (MLLMCU_ATTN_PROF=1): CTA 0 {entry, dependency
// wait passed, q rotated, key loop done, partial stored, ticket taken}, then the merging CTA of row 0 {merge start, output stored}
MLLMCU_API int mllmcu_decode_attn_prof_read(unsigned long long *host_out, int n_launches) {
    if (g_attn_prof_on != 1 || !host_out || n_launches <= 0) return 0;
    if (cudaDeviceSynchronize() != cudaSuccess) return 0;
    int n = n_launches < g_attn_prof_n ? n_launches : g_attn_prof_n;
    if (n > ATTN_PROF_RING) n = ATTN_PROF_RING;
    for (int i = 0; i < n; i++) {
        const int slot = (g_attn_prof_n - n + i) % ATTN_PROF_RING;
        if (cudaMemcpy(host_out + (size_t)i * 8, g_attn_prof + (size_t)slot * 8, 64, cudaMemcpyDeviceToHost) != cudaSuccess) return i;
    }
    return n;
}

// ------------------------------------------------------------------ greedy pick with token log
// ids_out[b] (fp32, the next step's input_ids) and tok_log[ctl[0] * rows + b] (int32): the token chosen
// after consuming position ctl[0].  First maximum wins (Generate.cpp:46-53).
__global__ void __launch_bounds__(1024) k_decode_argmax(const float *__restrict__ x, int n, float *__restrict__ ids_out, int *__restrict__ tok_log,
                                                        const int *__restrict__ ctl, int log_rows) {
    const float *xr = x + (size_t)blockIdx.x * n;
    float best = -INFINITY;
    int bi = 0x7fffffff;
    if ((n & 3) == 0 && (((uintptr_t)xr) & 15) == 0) {
        // 16-byte loads, eight in flight per thread: the scalar loop paid one L2 round trip per element and thread (18 us for 32 000 logits)
        const float4 *x4 = (const float4 *)xr;
        const int n4 = n >> 2;
        for (int i0 = threadIdx.x; i0 < n4; i0 += 8 * blockDim.x) {
            float4 v[8];
#pragma unroll
            for (int u = 0; u < 8; u++) {
                const int i = i0 + u * blockDim.x;
                v[u] = i < n4 ? x4[i] : make_float4(-INFINITY, -INFINITY, -INFINITY, -INFINITY);
            }
#pragma unroll
            for (int u = 0; u < 8; u++) { // ascending indices per thread, strict '>' keeps the first maximum; ties across threads resolve on the index below
                const int i = 4 * (i0 + u * blockDim.x);
                if (v[u].x > best) { best = v[u].x; bi = i; }
                if (v[u].y > best) { best = v[u].y; bi = i + 1; }
                if (v[u].z > best) { best = v[u].z; bi = i + 2; }
                if (v[u].w > best) { best = v[u].w; bi = i + 3; }
            }
        }
    } else {
        for (int i = threadIdx.x; i < n; i += blockDim.x) {
            const float v = xr[i];
            if (v > best || (v == best && i < bi)) { best = v; bi = i; }
        }
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
