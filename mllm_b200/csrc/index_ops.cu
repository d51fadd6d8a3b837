// index_ops.cu -- the index-producing / index-consuming tensor functions the VLM graphs use to splice image embeddings into the text
// prefill (models/llava/modeling_llava.hpp:128-131, models/qwen2_vl/modeling_qwen2_vl.hpp:386-389): Tensor::where, Tensor::index_put,
// Tensor::range.  All indices stay on the device; the host only reads one count (where's output shape is data dependent).
#include "common.cuh"
#include "mllmcu.h"

// ---- where: ordered compaction of the positions with x == value.  One CTA walks the tensor in chunks of its size; inside a chunk
// positions are ranked with ballots + a shared prefix over the warps, so the output order is the reference's visit order
// (CPUwhereFunction::execute, op/CPUWhereFunc.hpp:38-53: b, then s, then h, then d -- the memory order of a dense BSHD tensor).
__global__ void __launch_bounds__(1024) k_where_eq(const float *__restrict__ x, long long total, float value, int *__restrict__ lin, int *__restrict__ count) {
    __shared__ int warp_n[32];
    __shared__ int base;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid == 0) base = 0;
    __syncthreads();
    for (long long i0 = 0; i0 < total; i0 += blockDim.x) {
        const long long i = i0 + tid;
        const bool hit = i < total && x[i] == value;
        const unsigned m = __ballot_sync(0xffffffffu, hit);
        if (lane == 0) warp_n[warp] = __popc(m);
        __syncthreads();
        int before = 0;
        for (int w = 0; w < warp; w++) before += warp_n[w];
        if (hit) lin[base + before + __popc(m & ((1u << lane) - 1u))] = (int)i;
        __syncthreads();
        if (tid == 0) {
            int n = 0;
            for (int w = 0; w < (int)(blockDim.x >> 5); w++) n += warp_n[w];
            base += n;
        }
        __syncthreads();
    }
    if (tid == 0) *count = base;
}
MLLMCU_API int mllmcu_where_eq(const float *x, long long total, float value, int *lin_idx, int *count, void *stream) {
    if (total < 0 || total > 0x7fffffffLL) return -1;
    k_where_eq<<<1, 1024, 0, (cudaStream_t)stream>>>(x, total, value, lin_idx, count);
    CU_LAUNCH_CHECK();
    return 0;
}
// coordinate `axis` (0 b, 1 s, 2 h, 3 d in visit order, sizes n[4]) of each compacted position, as the float the reference stores
__global__ void k_where_unravel(const int *__restrict__ lin, int count, int n1, int n2, int n3, int axis, float *__restrict__ out) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= count) return;
    int v = lin[i];
    const int d = v % n3; v /= n3;
    const int h = v % n2; v /= n2;
    const int s = v % n1; v /= n1;
    out[i] = (float)(axis == 0 ? v : axis == 1 ? s : axis == 2 ? h : d);
}
MLLMCU_API int mllmcu_where_unravel(const int *lin_idx, int count, const int *n, int axis, float *out, void *stream) {
    if (count <= 0) return 0;
    if (axis < 0 || axis > 3) return -1;
    k_where_unravel<<<(count + 255) / 256, 256, 0, (cudaStream_t)stream>>>(lin_idx, count, n[1], n[2], n[3], axis, out);
    CU_LAUNCH_CHECK();
    return 0;
}

// ---- index_put, accumulate == false (op/CPUIndexPutFunc.hpp:81-88): dst row (int)idx[r] <- src row r
__global__ void k_index_put_rows(float *__restrict__ dst, long long dst_ld, const float *__restrict__ src, long long src_ld, const float *__restrict__ idx, int dim) {
    const int r = blockIdx.x;
    float *d = dst + (long long)(int)idx[r] * dst_ld;
    const float *s = src + (long long)r * src_ld;
    for (int i = threadIdx.x; i < dim; i += blockDim.x) d[i] = s[i];
}
MLLMCU_API int mllmcu_index_put_rows(float *dst, long long dst_ld, const float *src, long long src_ld, const float *idx, int n_rows, int dim, void *stream) {
    if (n_rows <= 0) return 0;
    k_index_put_rows<<<n_rows, 256, 0, (cudaStream_t)stream>>>(dst, dst_ld, src, src_ld, idx, dim);
    CU_LAUNCH_CHECK();
    return 0;
}
// ---- index_put, accumulate == true with one replaced position (op/CPUIndexPutFunc.hpp:89-118, replace_size == 1): the row at
// p = (int)idx[0] of dest [S][dim] is replaced by the R rows of src; out [S - 1 + R][dim] = dest[0:p] ++ src ++ dest[p+1:S]
__global__ void k_index_splice(const float *__restrict__ dest, const float *__restrict__ src, const float *__restrict__ idx, float *__restrict__ out, int R, int dim) {
    const int s = blockIdx.x, p = (int)idx[0];
    const float *from = s < p ? dest + (long long)s * dim : s < p + R ? src + (long long)(s - p) * dim : dest + (long long)(s - R + 1) * dim;
    float *to = out + (long long)s * dim;
    for (int i = threadIdx.x; i < dim; i += blockDim.x) to[i] = from[i];
}
MLLMCU_API int mllmcu_index_splice(const float *dest, const float *src, const float *idx, float *out, int S, int R, int dim, void *stream) {
    if (S <= 0 || R <= 0 || dim <= 0) return -1;
    k_index_splice<<<S - 1 + R, 256, 0, (cudaStream_t)stream>>>(dest, src, idx, out, R, dim);
    CU_LAUNCH_CHECK();
    return 0;
}

// ---- range (op/CPURangeFunc.hpp:33-41): out[i] = (float)(start + i)
__global__ void k_range(float *__restrict__ out, int start, int n) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) out[i] = (float)(start + i);
}
MLLMCU_API int mllmcu_range(float *out, int start, int n, void *stream) {
    if (n <= 0) return 0;
    k_range<<<(n + 255) / 256, 256, 0, (cudaStream_t)stream>>>(out, start, n);
    CU_LAUNCH_CHECK();
    return 0;
}
