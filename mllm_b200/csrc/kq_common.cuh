// kq_common.cuh -- pieces of the Q4_K super-block decode shared by kquants.cu and gemm_mma_q4k.cu
#pragma once
#include <stdint.h>

// get_scale_min_k4 (ggml/QuantizeQ4.cpp:177-184) for all eight sub-blocks of a super-block: s = the 12 scale bytes as three words;
// sc / mn receive the eight 6-bit values packed one per byte (sub-blocks 0-3 in [0], 4-7 in [1])
static __device__ __forceinline__ void q4k_scales8(uint32_t s0, uint32_t s1, uint32_t s2, uint32_t (&sc)[2], uint32_t (&mn)[2]) {
    sc[0] = s0 & 0x3F3F3F3Fu;                                           // q[j] & 63, j < 4
    mn[0] = s1 & 0x3F3F3F3Fu;                                           // q[j + 4] & 63
    sc[1] = (s2 & 0x0F0F0F0Fu) | ((s0 >> 2) & 0x30303030u);             // (q[j + 4] & 0xF) | ((q[j - 4] >> 6) << 4), j >= 4
    mn[1] = ((s2 >> 4) & 0x0F0F0F0Fu) | ((s1 >> 2) & 0x30303030u);      // (q[j + 4] >> 4) | ((q[j] >> 6) << 4)
}
static __device__ __forceinline__ int byte_of(const uint32_t (&v)[2], int j) { return (int)((v[j >> 2] >> (8 * (j & 3))) & 0xFFu); }
