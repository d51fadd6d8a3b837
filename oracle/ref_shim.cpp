// ref_shim.cpp -- C-ABI wrappers over the REFERENCE's own compiled sources.
//
// TEST INFRASTRUCTURE ONLY.  This file contains no arithmetic of its own: it
// only forwards to functions whose definitions are compiled from the sources
// where they lie under /root/reference (see oracle/Makefile; nothing from the
// reference is copied into this repo).  The resulting library,
// oracle/_ref/libmllm_ref.so, is used (a) to pin oracle.c, (b) to generate the
// golden fixtures under tests/golden/, (c) as bench.py's CPU baseline
// ("kind": "reference").  It is built only where /root/reference exists.
#include <cstdint>
#include <cstring>
#include <cmath>
#include <omp.h>

#include "Types.hpp"
#include "DataType.hpp"
#include "backends/cpu/third_party/ggml/QuantizeQ8.hpp"
#include "backends/cpu/third_party/ggml/QuantizeQ4.hpp"
#include "backends/cpu/third_party/ggml/VecDotQ4.hpp"
#include "backends/cpu/third_party/ggml/VecDotQ8.hpp"
#include "backends/cpu/third_party/ggml/VecDotQ6.hpp"
#include "backends/cpu/third_party/ggml/VecDotFP32.hpp"
#include "backends/cpu/third_party/ggml/QuantizeQ6.hpp"
#include "backends/cpu/third_party/ggml/Quantize.hpp"
#include "backends/cpu/third_party/ggml/GemmPack.hpp"
#include "backends/cpu/compute/ActivationFunction.hpp"
#include "backends/cpu/compute/FlashAttention2.hpp"

extern "C" {

// ggml/QuantizeQ8.cpp:57
void ref_quantize_row_q8_0(const float *x, void *y, int k) { quantize_row_q8_0(x, y, k); }
// ggml/QuantizeQ8.cpp:196
void ref_dequantize_row_q8_0(const void *x, float *y, int k) { dequantize_row_q8_0(x, y, k); }
// ggml/QuantizeQ4.cpp:70
void ref_quantize_row_q4_0(const float *x, void *y, int k) { quantize_row_q4_0(x, y, k); }
// ggml/QuantizeQ4.cpp:74
void ref_dequantize_row_q4_0(const void *x, float *y, int k) { dequantize_row_q4_0(x, y, k); }
// ggml/VecDotQ4.cpp:696
void ref_vec_dot_q4_0_q8_0(int n, float *s, const void *vx, const void *vy) { vec_dot_q4_0_q8_0(n, s, vx, vy); }
// ggml/VecDotQ8.cpp
void ref_vec_dot_q8_0_q8_0(int n, float *s, const void *vx, const void *vy) { vec_dot_q8_0_q8_0(n, s, vx, vy); }

uint16_t ref_fp32_to_fp16(float f) { return (uint16_t)MLLM_FP32_TO_FP16(f); }
float ref_fp16_to_fp32(uint16_t h) { return MLLM_FP16_TO_FP32(h); }

// compute/ActivationFunction.cpp:5
void ref_silu(const float *x, float *y, size_t n) { mllm::mllm_vec_silu_f32((int)n, y, x); }

// The loop nest of mat_mul's generic path (compute/Matmul.cpp:104-115 row
// quantisation, :223-276 vec_dot per (m,n) + bias + optional fp16 store),
// driving the reference's own quantize_row_q8_0 / vec_dot_* bodies.
int ref_linear(int wtype, const void *W, const float *x, const float *bias, void *y, int M, int N, int K, int ldx,
               int ldy, int out_fp16) {
    if (K % 32 || (wtype != 2 && wtype != 8)) return -1;
    const int nb = K / 32;
    const size_t wrow = (size_t)nb * (wtype == 2 ? sizeof(block_q4_0) : sizeof(block_q8_0));
    block_q8_0 *xq = (block_q8_0 *)malloc((size_t)M * nb * sizeof(block_q8_0));
#pragma omp parallel for
    for (int m = 0; m < M; m++) quantize_row_q8_0(x + (size_t)m * ldx, xq + (size_t)m * nb, K);
#pragma omp parallel for collapse(2)
    for (int m = 0; m < M; m++)
        for (int n = 0; n < N; n++) {
            float tmp = 0;
            const char *wr = (const char *)W + (size_t)n * wrow;
            if (wtype == 2) vec_dot_q4_0_q8_0(K, &tmp, wr, xq + (size_t)m * nb);
            else vec_dot_q8_0_q8_0(K, &tmp, wr, xq + (size_t)m * nb);
            if (out_fp16) {
                ((mllm_fp16_t *)y)[(size_t)m * ldy + n] = MLLM_FP32_TO_FP16(bias ? tmp + bias[n] : tmp);
            } else {
                float *o = (float *)y + (size_t)m * ldy + n;
                *o = tmp;
                if (bias) *o += bias[n];
            }
        }
    free(xq);
    return 0;
}

// compute/FlashAttention2.hpp:2236 with the arguments CPUFlashAttention2Func
// passes (op/CPUFlashAttention2Func.hpp:70-120): br = bc = (Sq >= 4 ? 4 : 1),
// high_precision_exp = true, fp16 K/V.  K/V must be contiguous [Sk,Hkv,D].
void ref_flash_attn_f16kv(const float *Q, const uint16_t *K, const uint16_t *V, float *O, int Sq, int Sk, int Hq,
                          int Hkv, int D, int causal, int threads) {
    int br = Sq >= 4 ? 4 : 1;
    if (threads > Hkv) threads = Hkv;
    if (threads < 1) threads = 1;
    while (Hq % threads) threads--;
    flash_attention_2_forward(Q, K, V, O, 1, Hq, Sq, Sk, D, causal != 0, false, threads, br, br, Hq, Hkv, true);
}

void ref_flash_attn_f32kv(const float *Q, const float *K, const float *V, float *O, int Sq, int Sk, int Hq, int Hkv,
                          int D, int causal, int threads) {
    int br = Sq >= 4 ? 4 : 1;
    if (threads > Hkv) threads = Hkv;
    if (threads < 1) threads = 1;
    while (Hq % threads) threads--;
    flash_attention_2_forward(Q, K, V, O, 1, Hq, Sq, Sk, D, causal != 0, true, threads, br, br, Hq, Hkv, true);
}

// ---- rows a4/a5/a7/a15/a17
// ggml/QuantizeQ8.cpp:262
void ref_quantize_row_q8_K(const float *x, void *y, int k) { quantize_row_q8_K(x, y, k); }
// ggml/QuantizeQ4.cpp:335 / :295
void ref_quantize_row_q4_K(const float *x, void *y, int k) { quantize_row_q4_K(x, y, k); }
void ref_dequantize_row_q4_K(const void *x, float *y, int k) { dequantize_row_q4_K((const block_q4_K *)x, y, k); }
// ggml/VecDotQ4.cpp:32
void ref_vec_dot_q4_K_q8_K(int n, float *s, const void *vx, const void *vy) { vec_dot_q4_K_q8_K(n, s, vx, vy); }
// ggml/QuantizeQ6.cpp:155 / :110, ggml/VecDotQ6.cpp:32
void ref_quantize_row_q6_K(const float *x, void *y, int k) { quantize_row_q6_K(x, y, k); }
void ref_dequantize_row_q6_K(const void *x, float *y, int k) { dequantize_row_q6_K((const block_q6_K *)x, y, k); }
void ref_vec_dot_q6_K_q8_K(int n, float *s, const void *vx, const void *vy) { vec_dot_q6_K_q8_K(n, s, vx, vy); }

// mat_mul generic path with a k-quant weight (compute/Matmul.cpp:104-115, :223-276): rows -> Q8_K, vec_dot per (m, n)
int ref_linear_kq(int wtype, const void *W, const float *x, const float *bias, float *y, int M, int N, int K, int ldx, int ldy) {
    if (K % QK_K || (wtype != 12 && wtype != 14)) return -1;
    const int nb = K / QK_K;
    const size_t wrow = (size_t)nb * (wtype == 12 ? sizeof(block_q4_K) : sizeof(block_q6_K));
    block_q8_K *xq = (block_q8_K *)calloc((size_t)M * nb, sizeof(block_q8_K));
    for (int m = 0; m < M; m++) quantize_row_q8_K(x + (size_t)m * ldx, xq + (size_t)m * nb, K);
#pragma omp parallel for collapse(2)
    for (int m = 0; m < M; m++)
        for (int n = 0; n < N; n++) {
            float tmp = 0;
            const char *wr = (const char *)W + (size_t)n * wrow;
            if (wtype == 12) vec_dot_q4_K_q8_K(K, &tmp, wr, xq + (size_t)m * nb);
            else vec_dot_q6_K_q8_K(K, &tmp, wr, xq + (size_t)m * nb);
            y[(size_t)m * ldy + n] = tmp;
            if (bias) y[(size_t)m * ldy + n] += bias[n];
        }
    free(xq);
    return 0;
}

// ggml/GemmPack.cpp:346 (the 4x4 interleave the quantizer writes for MLLM_TYPE_Q4_0_4_4) and :367 gemv
void ref_quantize_q4_0_4x4(const float *src, void *dst, int nrows, int k) {
    static const float dummy = 0.f; // non-null imatrix selects the 4x4 layout (GemmPack.cpp:348)
    quantize_q4_0_4x4(src, dst, nrows, k, &dummy);
}
// Matmul.cpp:158-217 decode shape: x row -> Q8_0, gemv over nc interleaved rows
void ref_gemv_q4_0_4x4(const void *Wp, const float *x, float *y, int N, int K) {
    block_q8_0 *xq = (block_q8_0 *)malloc((size_t)(K / 32) * sizeof(block_q8_0));
    quantize_row_q8_0(x, xq, K);
    gemv_q4_0_4x4_q8_0(K, y, 0, Wp, xq, 1, N, nullptr);
    free(xq);
}

// compute/ActivationFunction.cpp:29 + the row loop of CPUSoftMax::execute (op/CPUSoftMax.cpp:41-62)
void ref_softmax(const float *x, float *y, int outer, int S, int n, int causal, int old_dim) {
    memset(y, 0, sizeof(float) * (size_t)outer * S * n);
    for (int o = 0; o < outer; o++)
        for (int s = 0; s < S; s++) {
            const float *xr = x + ((size_t)o * S + s) * n;
            float *dp = y + ((size_t)o * S + s) * n;
            int masked = (causal && S > 1) ? s + 1 + old_dim : n;
            if (masked > n) masked = n;
            float max = -INFINITY;
            for (int j = 0; j < masked; ++j) max = MAX(max, xr[j]);
            float sum = mllm::mllm_vec_soft_max_f32(masked, dp, xr, max);
            sum = 1.0 / sum;
            for (int j = 0; j < masked; ++j) dp[j] *= sum; // vec_scale_f32
        }
}

// ggml/Quantize.hpp:74-133: the fp16 lookup tables and their use
void ref_gelu(int kind, const float *x, float *y, int n) {
    if (kind == 0) { init_table_gelu_f16(); mllm_vec_gelu_f32(n, y, x); }
    else { init_table_gelu_quick_f16(); mllm_vec_gelu_quick_f32(n, y, x); }
}
void ref_gelu_table(int kind, uint16_t *out) {
    if (kind == 0) { init_table_gelu_f16(); memcpy(out, mllm_table_gelu_f16, sizeof(mllm_table_gelu_f16)); }
    else { init_table_gelu_quick_f16(); memcpy(out, mllm_table_gelu_quick_f16, sizeof(mllm_table_gelu_quick_f16)); }
}

// conv2d_fp32_VALID (compute/Convolution.cpp:35-90) needs Tensor objects; this drives the same vec_dot_fp32
// (ggml/VecDotFP32.cpp) over the same (c, kh, kw)-ordered receptive-field rows.
void ref_conv2d_valid(const float *x, const float *w, const float *bias, float *y, int B, int C, int H, int W, int O,
                      int kh, int kw, int sh, int sw) {
    const int oh = (H - kh) / sh + 1, ow = (W - kw) / sw + 1, KK = C * kh * kw;
    float *row = (float *)malloc(sizeof(float) * KK);
    for (int b = 0; b < B; b++)
        for (int i = 0; i < oh; i++)
            for (int j = 0; j < ow; j++) {
                for (int c = 0; c < C; c++)
                    for (int p = 0; p < kh; p++)
                        for (int q = 0; q < kw; q++) row[(c * kh + p) * kw + q] = x[(((size_t)b * C + c) * H + i * sh + p) * W + j * sw + q];
                for (int o = 0; o < O; o++) {
                    float v = 0;
                    vec_dot_fp32(KK, &v, w + (size_t)o * KK, row);
                    if (bias) v += bias[o];
                    y[(((size_t)b * O + o) * oh + i) * ow + j] = v;
                }
            }
    free(row);
}

int ref_max_threads() { return omp_get_max_threads(); }
}
