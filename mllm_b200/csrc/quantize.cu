// quantize.cu -- Q8_0 activation-row quantisation.
// Reference: quantize_row_q8_0, AVX2 body (ggml/QuantizeQ8.cpp:106-160), called per
// activation row by mat_mul (compute/Matmul.cpp:104-115).  One warp per 32-float
// block, one lane per element: amax by shuffle, then the exact reference sequence.
// HBM-bound elementwise work: reads 4 B/elt, writes ~1.06 B/elt.
#include "common.cuh"
#include "quantize.cuh"

// native block_q8_0 output (public API; bytes identical to the reference's)
__global__ void k_quantize_q8_0_native(const float *__restrict__ x, block_q8_0 *__restrict__ y, int rows, int nb, int ldx) {
    const int lane = threadIdx.x & 31;
    const long long wid = (long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (wid >= (long long)rows * nb) return;
    const int r = (int)(wid / nb), b = (int)(wid % nb);
    const float v = x[(size_t)r * ldx + (size_t)b * QK + lane];
    const float amax = warp_max(fabsf(v));
    block_q8_0 *o = y + (size_t)r * nb + b;
    o->qs[lane] = (int8_t)q8_0_quant(v, amax);
    if (lane == 0) o->d = q8_0_scale(amax);
}

// internal layout for the Linear kernels: qs[rows][k] int8 (16B-aligned rows),
// d[rows][nb] fp32 (= fp32(fp16 scale), what MLLM_FP16_TO_FP32(y[i].d) yields).
__global__ void k_quantize_q8_0_split(const float *__restrict__ x, int8_t *__restrict__ qs, float *__restrict__ d, int rows, int nb, int ldx) {
    const int lane = threadIdx.x & 31;
    const long long wid = (long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (wid >= (long long)rows * nb) return;
    const int r = (int)(wid / nb), b = (int)(wid % nb);
    const float v = x[(size_t)r * ldx + (size_t)b * QK + lane];
    const float amax = warp_max(fabsf(v));
    qs[((size_t)r * nb + b) * QK + lane] = (int8_t)q8_0_quant(v, amax);
    if (lane == 0) d[(size_t)r * nb + b] = __half2float(q8_0_scale(amax));
}

int launch_quantize_q8_0_split(const float *x, int8_t *qs, float *d, int rows, int k, int ldx, cudaStream_t st) {
    const int nb = k / QK;
    const long long warps = (long long)rows * nb;
    const int wpb = 8;
    k_quantize_q8_0_split<<<(unsigned)((warps + wpb - 1) / wpb), wpb * 32, 0, st>>>(x, qs, d, rows, nb, ldx);
    CU_LAUNCH_CHECK();
    return 0;
}

MLLMCU_API int mllmcu_quantize_q8_0(const float *x, void *y, int rows, int k, int ldx, void *stream) {
    if (rows <= 0 || k <= 0 || k % QK) return -1;
    const int nb = k / QK;
    const long long warps = (long long)rows * nb;
    const int wpb = 8;
    k_quantize_q8_0_native<<<(unsigned)((warps + wpb - 1) / wpb), wpb * 32, 0, (cudaStream_t)stream>>>(x, (block_q8_0 *)y, rows, nb, ldx);
    CU_LAUNCH_CHECK();
    return 0;
}
