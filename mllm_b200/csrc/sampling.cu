// sampling.cu -- top-k and top-p sampling on the device: _LlmTextGenerateTopkSamplingMethod::generate (Generate.cpp:45-91) and
// _LlmTextGenerateToppSamplingMethod::generate (Generate.cpp:93-143), so that a sampled token, like the greedy one
// (k_decode_argmax), never makes the vocabulary-sized logits row cross PCIe.
//
// Reference arithmetic kept: candidates in descending order of value (std::partial_sort / std::sort; equal values are taken lowest
// index first here, the reference leaves their order unspecified); softmax over the candidates with temperature,
//   e_i = (float) exp(((double)x_i - (double)max) / T),  s = sum e_i in double,  p_i = (float)(e_i / s),  then p_i /= (float) sum p_i
// (the "renormalise" of Generate.cpp:84-87), and the draw of std::discrete_distribution: weights / their double sum, running sums,
// first bucket whose running sum exceeds u.  The reference draws u from a std::mt19937 seeded by std::random_device, i.e. not
// reproducibly; here the caller passes u in [0, 1) (one per row), so the same u gives the same token on every run.
// top-p expects probabilities (the reference throws when the largest input exceeds 1): candidates are taken until their
// float running sum reaches p.
#include "common.cuh"
#include "mllmcu.h"

#define SAMPLE_MAX_CAND 1024

// block-wide argmax of v[0..n) (first index wins among equal values); all threads return the result
static __device__ int block_argmax(const float *__restrict__ v, int n, float *s_val, int *s_idx, float *best_out) {
    float bv = -INFINITY;
    int bi = 0x7fffffff;
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
        const float x = v[i];
        if (x > bv || (x == bv && i < bi)) { bv = x; bi = i; }
    }
    for (int o = 16; o > 0; o >>= 1) {
        const float ov = __shfl_xor_sync(0xffffffffu, bv, o);
        const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
        if (ov > bv || (ov == bv && oi < bi)) { bv = ov; bi = oi; }
    }
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (lane == 0) { s_val[warp] = bv; s_idx[warp] = bi; }
    __syncthreads();
    if (warp == 0) {
        bv = lane < (int)(blockDim.x >> 5) ? s_val[lane] : -INFINITY;
        bi = lane < (int)(blockDim.x >> 5) ? s_idx[lane] : 0x7fffffff;
        for (int o = 16; o > 0; o >>= 1) {
            const float ov = __shfl_xor_sync(0xffffffffu, bv, o);
            const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
            if (ov > bv || (ov == bv && oi < bi)) { bv = ov; bi = oi; }
        }
        if (lane == 0) { s_val[0] = bv; s_idx[0] = bi; }
    }
    __syncthreads();
    const int r = s_idx[0];
    *best_out = s_val[0];
    __syncthreads();
    return r;
}

// one CTA per row.  mode 0: top-k (k candidates);  mode 1: top-p (candidates until their sum >= p_or_k).
__global__ void __launch_bounds__(1024) k_sample(const float *__restrict__ x, int n, long long ldx, int mode, float p_or_k, float temperature, const float *__restrict__ u,
                                                 float *__restrict__ ws, float *__restrict__ ids_out, float *__restrict__ cand_p, int *__restrict__ cand_idx,
                                                 int *__restrict__ n_cand_out, int cand_cap) {
    __shared__ float s_val[32];
    __shared__ int s_idx[32];
    __shared__ float c_val[SAMPLE_MAX_CAND];
    __shared__ int c_idx[SAMPLE_MAX_CAND];
    __shared__ int s_n;
    const int row = blockIdx.x;
    const float *xr = x + (long long)row * ldx;
    float *w = ws + (long long)row * n;
    for (int i = threadIdx.x; i < n; i += blockDim.x) w[i] = xr[i];
    __syncthreads();
    const int k = mode == 0 ? min((int)p_or_k, min(n, SAMPLE_MAX_CAND)) : min(n, SAMPLE_MAX_CAND);
    float cum = 0.f; // top-p: float running sum, Generate.cpp:108-115
    int cnt = 0;
    while (cnt < k) {
        float bv;
        const int bi = block_argmax(w, n, s_val, s_idx, &bv);
        if (threadIdx.x == 0) { c_val[cnt] = bv; c_idx[cnt] = bi; w[bi] = -INFINITY; }
        cnt++;
        __syncthreads();
        if (mode == 1) {
            cum += bv;
            if (!(cum < p_or_k)) break;
        }
    }
    if (threadIdx.x == 0) {
        int pick = 0;
        if (cnt > 1) {
            // softmax with temperature (Generate.cpp:69-87 / 121-139), then std::discrete_distribution's draw
            const double mx = (double)c_val[0];
            double sum_exp = 0.0;
            for (int i = 0; i < cnt; i++) {
                const float e = (float)exp(((double)c_val[i] - mx) / (double)temperature);
                c_val[i] = e;
                sum_exp += (double)e;
            }
            double s2 = 0.0;
            for (int i = 0; i < cnt; i++) { c_val[i] = (float)((double)c_val[i] / sum_exp); s2 += (double)c_val[i]; }
            const float s2f = (float)s2;
            double tot = 0.0;
            for (int i = 0; i < cnt; i++) { c_val[i] = c_val[i] / s2f; tot += (double)c_val[i]; }
            const double uu = (double)u[row];
            double run = 0.0;
            pick = cnt - 1;
            for (int i = 0; i < cnt; i++) {
                run += (double)c_val[i] / tot;
                if (uu < run) { pick = i; break; }
            }
        } else {
            c_val[0] = 1.f;
        }
        ids_out[row] = (float)c_idx[pick];
        // top-p that ran into the candidate limit before reaching p: reported as -count (the sample is from the truncated set)
        if (n_cand_out) n_cand_out[row] = (mode == 1 && cum < p_or_k && cnt < n) ? -cnt : cnt;
        s_n = cnt;
    }
    __syncthreads();
    if (cand_p && cand_idx)
        for (int i = threadIdx.x; i < min(s_n, cand_cap); i += blockDim.x) {
            cand_p[(long long)row * cand_cap + i] = c_val[i];
            cand_idx[(long long)row * cand_cap + i] = c_idx[i];
        }
}

MLLMCU_API int mllmcu_sample(int mode, const float *x, int rows, int n, long long ldx, float p_or_k, float temperature, const float *u, float *ws, float *ids_out,
                             float *cand_p, int *cand_idx, int *n_cand, int cand_cap, void *stream) {
    if (rows <= 0 || n <= 0 || (mode != 0 && mode != 1) || !(temperature > 0.f) || !ws || !u) return -1;
    if (mode == 0 && (p_or_k < 2.f || p_or_k > (float)SAMPLE_MAX_CAND)) return -1; // k = 0 / 1 is the greedy pick (Generate.cpp:50-54): mllmcu_argmax_ids
    k_sample<<<rows, 1024, 0, (cudaStream_t)stream>>>(x, n, ldx, mode, p_or_k, temperature, u, ws, ids_out, cand_p, cand_idx, n_cand, cand_cap);
    CU_LAUNCH_CHECK();
    return 0;
}
