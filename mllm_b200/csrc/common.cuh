// common.cuh -- shared device helpers for the mllm CUDA backend kernels (sm_100a).
#pragma once
#include <cuda_runtime.h>
#include <cuda_fp16.h>
#include <stdint.h>

#define MLLMCU_API extern "C" __attribute__((visibility("default")))

#define QK 32
// Native block layouts, byte-identical to mllm/DataType.hpp:75-78 (block_q4_0)
// and :137-140 (block_q8_0).  Kept packed: rows of these are uploaded as-is.
struct __align__(2) block_q4_0 { __half d; uint8_t qs[16]; };
struct __align__(2) block_q8_0 { __half d; int8_t qs[32]; };
static_assert(sizeof(block_q4_0) == 18, "q4_0 block");
static_assert(sizeof(block_q8_0) == 34, "q8_0 block");

#define CU_TRY(x)                                   \
    do {                                            \
        cudaError_t e__ = (x);                      \
        if (e__ != cudaSuccess) return (int)e__;    \
    } while (0)
#define CU_LAUNCH_CHECK() CU_TRY(cudaGetLastError())

// Multiprocessor count of the current device (grid sizing: persistent kernels launch one CTA per SM, split-K heuristics aim at
// a multiple of it); queried once per device, 148 on a B200.
static inline int cu_sm_count() {
    static int cached[64] = {};
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
    if (cached[dev] == 0) {
        int n = 0;
        if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
        cached[dev] = n;
    }
    return cached[dev];
}

// RoPE rotation of the pair (a, b) by (sin, cos): the reference writes `a*cos - b*sin` and `a*sin + b*cos` in scalar C++
// (op/CPURoPE.cpp:241-242, CPUMultimodalRoPE.cpp:171-172, CPUVisionRoPEFunc.hpp:44-45); compiled for an FMA target gcc contracts
// the FIRST product of each into the add -- fma(a, cos, -(b*sin)), fma(a, sin, b*cos) -- which is what its users get and what the
// compiled reference Op classes return (tests/test_oracle_refops.py pins it).  Written out so nvcc cannot choose differently.
static __device__ __forceinline__ float rope_lo(float a, float b, float sn, float cs) { return __fmaf_rn(a, cs, -__fmul_rn(b, sn)); }
static __device__ __forceinline__ float rope_hi(float a, float b, float sn, float cs) { return __fmaf_rn(a, sn, __fmul_rn(b, cs)); }

static __device__ __forceinline__ float warp_max(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}
static __device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
static __device__ __forceinline__ double warp_sum_d(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
static __device__ __forceinline__ int warp_sum_i(int v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// Q8_0 quantisation of one value given the block's amax; bit-identical to the
// AVX2 body of quantize_row_q8_0 (ggml/QuantizeQ8.cpp:128-150): id = 127/amax
// is its own IEEE division, the product is rounded once (no FMA contraction)
// and converted with round-half-even.
static __device__ __forceinline__ int q8_0_quant(float x, float amax) {
    const float id = (amax != 0.0f) ? __fdiv_rn(127.f, amax) : 0.0f;
    return __float2int_rn(__fmul_rn(x, id));
}
static __device__ __forceinline__ __half q8_0_scale(float amax) { return __float2half_rn(__fdiv_rn(amax, 127.f)); }

// 128-bit streaming load that does not allocate in L1 (weights are read once).
static __device__ __forceinline__ uint4 ld_stream_u4(const void *p) {
    uint4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
    return r;
}

// mllm_v_expf (compute/ActivationFunction.hpp:96-141) per element: same fused / unfused operations as the AVX2 and
// AVX-512 vector bodies, so results are bit-identical to theirs.
static __device__ __forceinline__ float v_expf(float x) {
    const float r = 0x1.8p23f;
    const float z = __fmaf_rn(x, 0x1.715476p+0f, r);
    const float n = __fsub_rn(z, r);
    const float b = __fmaf_rn(-n, 0x1.7f7d1cp-20f, __fmaf_rn(-n, 0x1.62e4p-1f, x));
    const uint32_t e = __float_as_uint(z) << 23;
    const float k = __uint_as_float(e + __float_as_uint(1.0f));
    const bool c = fabsf(n) > 126.0f;
    const float u = __fmul_rn(b, b);
    const float j = __fmaf_rn(__fmaf_rn(__fmaf_rn(0x1.0e4020p-7f, b, 0x1.573e2ep-5f), u, __fmaf_rn(0x1.555e66p-3f, b, 0x1.fffdb6p-2f)), u,
                              __fmul_rn(0x1.ffffecp-1f, b));
    if (!c) return __fmaf_rn(j, k, k);
    const uint32_t g = (n <= 0.0f) ? 0x82000000u : 0u;
    const float s1 = __uint_as_float(g + 0x7f000000u);
    const float s2 = __uint_as_float(e - g);
    if (fabsf(n) > 192.0f) return __fmul_rn(s1, s1);
    return __fmul_rn(__fmaf_rn(s2, j, s2), s1);
}
