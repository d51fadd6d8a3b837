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

int ref_max_threads() { return omp_get_max_threads(); }
}
