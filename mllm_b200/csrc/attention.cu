// attention.cu -- flash-style attention over the fp16 KV cache.
// Reference: CPUFlashAttention2Func::execute -> flash_attention_2_forward
// (op/CPUFlashAttention2Func.hpp:49-125; compute/FlashAttention2.hpp:1212-2284,
// fp32 Q, fp16 K/V, fp32 accumulate): o = softmax(q.k^T / sqrt(D) + causal) v,
// online softmax, GQA by kv_head = q_head / (Hq/Hkv); key c visible to query r
// iff c <= r + (Sk - Sq).
//
// Note on a reference defect: the fp16-KV prefill body computes its leftover
// key tile as `seq_size_k % Tc` (FlashAttention2.hpp:1274) instead of `% Bc`,
// so for Sk % 4 != 0 it drops/mis-sizes the last keys.  This kernel implements
// the intended causal attention for every Sk; parity tests against the compiled
// reference therefore use Sk % 4 == 0 for prefill (decode is unaffected).
//
// Layouts: q,o [B][Sq][Hq][D] fp32;  k,v fp16 at b*kv_bstride + s*Hkv*D + h*D + d.
//
// k_attn_splitk: one warp per (b, q row, head, key split); lanes split D
// (D/32 contiguous elements each, so a key row is one coalesced 2*D-byte read);
// per-split (m, l, acc[D]) partials go to the workspace when nsplit > 1 and are
// merged by k_attn_merge.  Decode is HBM-bound on the KV read
// (2 * Sk * Hkv * D * 2 bytes per sequence).
#include "common.cuh"
#include <algorithm>
#include <stdlib.h>

static __device__ __forceinline__ float kv_to_f32(__half h) { return __half2float(h); }
static __device__ __forceinline__ float kv_to_f32(float f) { return f; }

// KT = __half (the KV cache) or float (ViT blocks, whose K/V never enter a cache: models/vit/modeling_vit.hpp:42-45,
// flash_attention_2_forward's fp32 variant, compute/FlashAttention2.hpp:87-1210).  D need not be a multiple of 32
// (Qwen2-VL's ViT has D = 80): then D/4 lanes carry 4 elements each and the rest idle.
template <int D, typename KT>
__global__ void __launch_bounds__(128) k_attn_splitk(const float *__restrict__ q, const KT *__restrict__ k,
                                                     const KT *__restrict__ v, float *__restrict__ o, float *__restrict__ ws,
                                                     int B, int Sq, int Sk, int Hq, int Hkv, int causal, long long kvb,
                                                     int nsplit, int chunk) {
    constexpr int E = (D % 32 == 0) ? D / 32 : 4;
    constexpr int ACT = D / E; // lanes that hold data
    const int lane = threadIdx.x & 31;
    const bool live = lane < ACT;
    const int lo = live ? lane * E : 0;
    const long long wid = (long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    const long long total = (long long)B * Sq * Hq * nsplit;
    if (wid >= total) return;
    const int sp = (int)(wid % nsplit);
    const int h = (int)((wid / nsplit) % Hq);
    const int r = (int)((wid / ((long long)nsplit * Hq)) % Sq);
    const int b = (int)(wid / ((long long)nsplit * Hq * Sq));
    const int kvh = h / (Hq / Hkv);
    const float scale = __fdiv_rn(1.0f, __fsqrt_rn((float)D));
    const int last = causal ? min(Sk - 1, r + (Sk - Sq)) : (Sk - 1);
    const int c_begin = sp * chunk;
    const int c_end = min(last + 1, c_begin + chunk);

    float qv[E];
    const float *qp = q + (((size_t)b * Sq + r) * Hq + h) * D + lo;
#pragma unroll
    for (int e = 0; e < E; e++) qv[e] = live ? qp[e] : 0.f;
    float acc[E];
#pragma unroll
    for (int e = 0; e < E; e++) acc[e] = 0.f;
    float m = -INFINITY, l = 0.f;
    const KT *kb = k + (size_t)b * kvb + (size_t)kvh * D + lo;
    const KT *vb = v + (size_t)b * kvb + (size_t)kvh * D + lo;
    const size_t rs = (size_t)Hkv * D;
    for (int c = c_begin; c < c_end; c++) {
        float kf[E], vf[E];
        if (E == 4 && sizeof(KT) == 2) {
            const uint2 kk = *(const uint2 *)(kb + (size_t)c * rs);
            const uint2 vv = *(const uint2 *)(vb + (size_t)c * rs);
            const __half2 *k2 = (const __half2 *)&kk, *v2 = (const __half2 *)&vv;
            float2 t;
            t = __half22float2(k2[0]); kf[0] = t.x; kf[1] = t.y;
            t = __half22float2(k2[1]); kf[2] = t.x; kf[3] = t.y;
            t = __half22float2(v2[0]); vf[0] = t.x; vf[1] = t.y;
            t = __half22float2(v2[1]); vf[2] = t.x; vf[3] = t.y;
        } else {
#pragma unroll
            for (int e = 0; e < E; e++) { kf[e] = kv_to_f32(kb[(size_t)c * rs + e]); vf[e] = kv_to_f32(vb[(size_t)c * rs + e]); }
        }
        float s = 0.f;
#pragma unroll
        for (int e = 0; e < E; e++) s = fmaf(qv[e], kf[e], s);
        s = warp_sum(s);
        const float m_new = fmaxf(m, s);
        const float sc = expf((m - m_new) * scale); // m=-inf first time -> 0
        const float p = expf((s - m_new) * scale);
        l = l * sc + p;
#pragma unroll
        for (int e = 0; e < E; e++) acc[e] = fmaf(p, vf[e], acc[e] * sc);
        m = m_new;
    }
    if (nsplit == 1) {
        float *op = o + (((size_t)b * Sq + r) * Hq + h) * D + lo;
        const float inv = 1.0f / l;
        if (live) {
#pragma unroll
            for (int e = 0; e < E; e++) op[e] = acc[e] * inv;
        }
    } else {
        float *wp = ws + (size_t)wid * (D + 2);
        if (lane == 0) { wp[0] = m; wp[1] = l; }
        if (live) {
#pragma unroll
            for (int e = 0; e < E; e++) wp[2 + lo + e] = acc[e];
        }
    }
}

template <int D>
__global__ void k_attn_merge(const float *__restrict__ ws, float *__restrict__ o, long long rows, int nsplit) {
    constexpr int E = (D % 32 == 0) ? D / 32 : 4;
    constexpr int ACT = D / E;
    const int lane = threadIdx.x & 31;
    const bool live = lane < ACT;
    const int lo = live ? lane * E : 0;
    const long long row = (long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (row >= rows) return;
    const float scale = __fdiv_rn(1.0f, __fsqrt_rn((float)D));
    float m = -INFINITY;
    for (int s = 0; s < nsplit; s++) m = fmaxf(m, ws[((size_t)row * nsplit + s) * (D + 2)]);
    float l = 0.f, acc[E];
#pragma unroll
    for (int e = 0; e < E; e++) acc[e] = 0.f;
    for (int s = 0; s < nsplit; s++) {
        const float *wp = ws + ((size_t)row * nsplit + s) * (D + 2);
        const float ms = wp[0];
        if (ms == -INFINITY) continue; // empty split (beyond the causal limit)
        const float f = expf((ms - m) * scale);
        l = fmaf(wp[1], f, l);
#pragma unroll
        for (int e = 0; e < E; e++) acc[e] = fmaf(wp[2 + lo + e], f, acc[e]);
    }
    const float inv = 1.0f / l;
    float *op = o + (size_t)row * D + lo;
    if (live) {
#pragma unroll
        for (int e = 0; e < E; e++) op[e] = acc[e] * inv;
    }
}

static int pick_nsplit(int B, int Sq, int Hq, int Sk) {
    // enough warps to cover every SM x 16 warps; a split should keep >= 32 keys
    const long long rows = (long long)B * Sq * Hq;
    long long want = ((long long)cu_sm_count() * 16 + rows - 1) / rows;
    long long maxs = (Sk + 31) / 32;
    if (want > maxs) want = maxs;
    if (want < 1) want = 1;
    if (want > 64) want = 64;
    return (int)want;
}

MLLMCU_API size_t mllmcu_flash_attn_workspace_bytes(int B, int Sq, int Hq, int D) {
    // worst case nsplit = 64 only happens when rows are few; bound rows*nsplit
    const long long rows = (long long)B * Sq * Hq;
    long long ns = ((long long)cu_sm_count() * 16 + rows - 1) / rows;
    if (ns > 64) ns = 64;
    if (ns < 1) ns = 1;
    return (size_t)rows * ns * (D + 2) * sizeof(float) + 256;
}

template <int D, typename KT>
static int launch_attn(const float *q, const KT *k, const KT *v, float *o, int B, int Sq, int Sk, int Hq, int Hkv,
                       int causal, long long kvb, float *ws, size_t ws_bytes, cudaStream_t st) {
    int nsplit = pick_nsplit(B, Sq, Hq, Sk);
    const long long rows = (long long)B * Sq * Hq;
    if (nsplit > 1 && (ws == nullptr || ws_bytes < (size_t)rows * nsplit * (D + 2) * sizeof(float))) nsplit = 1;
    const int chunk = (Sk + nsplit - 1) / nsplit;
    const long long warps = rows * nsplit;
    k_attn_splitk<D, KT><<<(unsigned)((warps + 3) / 4), 128, 0, st>>>(q, k, v, o, ws, B, Sq, Sk, Hq, Hkv, causal, kvb, nsplit, chunk);
    CU_LAUNCH_CHECK();
    if (nsplit > 1) {
        k_attn_merge<D><<<(unsigned)((rows + 3) / 4), 128, 0, st>>>(ws, o, rows, nsplit);
        CU_LAUNCH_CHECK();
    }
    return 0;
}

// attention_tc.cu: tensor-core prefill kernels; return 1 when the shape is not theirs
int attn_prefill_tc_f16(const float *q, const void *k, const void *v, float *o, int B, int Sq, int Sk, int Hq, int Hkv, int D, int causal,
                        long long kvb, cudaStream_t st);
size_t attn_tc_f32_planes_bytes(int B, int Sq, int Sk, int Hkv, int D);
int attn_prefill_tc_f32(const float *q, const void *k, const void *v, float *o, int B, int Sq, int Sk, int Hq, int Hkv, int D, int causal,
                        long long kvb, void *ws, size_t ws_bytes, cudaStream_t st);
static bool use_tc() { // MLLM_CUDA_ATTN_TC=0 keeps every shape on the CUDA-core kernel (A/B measurements)
    const char *e = getenv("MLLM_CUDA_ATTN_TC"); // read per call so one process can compare both
    return !(e && e[0] == '0');
}

MLLMCU_API int mllmcu_flash_attn_f16kv(const float *q, const void *k, const void *v, float *o, int B, int Sq, int Sk, int Hq,
                                       int Hkv, int D, int causal, long long kv_bstride, void *ws, size_t ws_bytes, void *stream) {
    if (B <= 0 || Sq <= 0 || Sk <= 0 || Hq <= 0 || Hkv <= 0 || Hq % Hkv) return -1;
    if (causal && Sq > Sk) return -1;
    cudaStream_t st = (cudaStream_t)stream;
    if (use_tc()) {
        const int rc = attn_prefill_tc_f16(q, k, v, o, B, Sq, Sk, Hq, Hkv, D, causal, kv_bstride, st);
        if (rc != 1) return rc;
    }
    switch (D) {
    case 64: return launch_attn<64, __half>(q, (const __half *)k, (const __half *)v, o, B, Sq, Sk, Hq, Hkv, causal, kv_bstride, (float *)ws, ws_bytes, st);
    case 128: return launch_attn<128, __half>(q, (const __half *)k, (const __half *)v, o, B, Sq, Sk, Hq, Hkv, causal, kv_bstride, (float *)ws, ws_bytes, st);
    case 96: return launch_attn<96, __half>(q, (const __half *)k, (const __half *)v, o, B, Sq, Sk, Hq, Hkv, causal, kv_bstride, (float *)ws, ws_bytes, st);
    case 32: return launch_attn<32, __half>(q, (const __half *)k, (const __half *)v, o, B, Sq, Sk, Hq, Hkv, causal, kv_bstride, (float *)ws, ws_bytes, st);
    default: return -2;
    }
}

// fp32 K/V (no cache): the attention of the vision towers -- CLIP ViT-L/14 (D 64), Qwen2-VL ViT (D 80) -- and of any graph
// that calls flash_attention_2_forward on fp32 tensors.  k,v: element (b,s,h,d) at b*kv_bstride + s*Hkv*D + h*D + d.
MLLMCU_API size_t mllmcu_flash_attn_f32kv_workspace_bytes(int B, int Sq, int Sk, int Hq, int Hkv, int D) {
    return std::max(mllmcu_flash_attn_workspace_bytes(B, Sq, Hq, D), attn_tc_f32_planes_bytes(B, Sq, Sk, Hkv, D));
}

MLLMCU_API int mllmcu_flash_attn_f32kv(const float *q, const float *k, const float *v, float *o, int B, int Sq, int Sk, int Hq,
                                       int Hkv, int D, int causal, long long kv_bstride, void *ws, size_t ws_bytes, void *stream) {
    if (B <= 0 || Sq <= 0 || Sk <= 0 || Hq <= 0 || Hkv <= 0 || Hq % Hkv) return -1;
    if (causal && Sq > Sk) return -1;
    cudaStream_t st = (cudaStream_t)stream;
    if (use_tc()) {
        const int rc = attn_prefill_tc_f32(q, k, v, o, B, Sq, Sk, Hq, Hkv, D, causal, kv_bstride, ws, ws_bytes, st);
        if (rc != 1) return rc;
    }
    switch (D) {
    case 64: return launch_attn<64, float>(q, k, v, o, B, Sq, Sk, Hq, Hkv, causal, kv_bstride, (float *)ws, ws_bytes, st);
    case 80: return launch_attn<80, float>(q, k, v, o, B, Sq, Sk, Hq, Hkv, causal, kv_bstride, (float *)ws, ws_bytes, st);
    case 128: return launch_attn<128, float>(q, k, v, o, B, Sq, Sk, Hq, Hkv, causal, kv_bstride, (float *)ws, ws_bytes, st);
    case 32: return launch_attn<32, float>(q, k, v, o, B, Sq, Sk, Hq, Hkv, causal, kv_bstride, (float *)ws, ws_bytes, st);
    default: return -2;
    }
}
