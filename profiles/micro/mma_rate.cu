// mma_rate.cu -- how fast does tcgen05.mma (kind::f16, M=128, N=128, K=16, cta_group::1) issue back to back with the
// K-major NO-SWIZZLE operand layout gemm_tc.cu uses?  One CTA per SM issues `iters` x `per` MMAs into one accumulator from fixed
// shared-memory operands (garbage values) and reports cycles per MMA.  mode 0: A and B from shared memory; 1: A from TMEM;
// 3: whole warp runs the loop, elect.sync lane issues; 4: as 3 with N = 256; 5: as 3 with commit + wait per group
// 2: as 0 but a commit + mbarrier wait after every `per` MMAs (the pipeline's hand-off granularity).
// build: nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -o mma_rate mma_rate.cu ; run: ./mma_rate
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
static __device__ __forceinline__ uint32_t su32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
static __device__ __forceinline__ uint64_t desc(uint32_t saddr, uint32_t lbo, uint32_t sbo) {
    return (uint64_t)((saddr >> 4) & 0x3FFF) | ((uint64_t)((lbo >> 4) & 0x3FFF) << 16) | ((uint64_t)((sbo >> 4) & 0x3FFF) << 32) | (1ull << 46);
}
#define IDESC ((1u << 4) | ((uint32_t)(128 >> 3) << 17) | ((uint32_t)(128 >> 4) << 24))
__global__ void __launch_bounds__(128, 1) k(int mode, int iters, int per, long long *out) {
    extern __shared__ __align__(1024) uint8_t smem[];
    __shared__ uint64_t bar;
    __shared__ uint32_t slot;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int i = threadIdx.x; i < 65536 / 4; i += 128) ((uint32_t *)smem)[i] = 0x3c003c00u; // 1.0 halves
    if (warp == 0) {
        if (lane == 0) { asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(su32(&bar))); asm volatile("fence.mbarrier_init.release.cluster;"); }
        __syncwarp();
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(su32(&slot)), "r"(512u) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tm = slot;
    if (mode >= 3 && warp == 0) { // whole warp runs the loop; one elected lane issues (CUTLASS style)
        const uint32_t a = su32(smem), b = su32(smem + 32768);
        uint32_t phase = 0;
        uint32_t el;
        asm volatile("{\n.reg .b32 rx;\n.reg .pred px;\nelect.sync rx|px, 0xffffffff;\nselp.b32 %0, 1, 0, px;\n}" : "=r"(el));
        long long t0 = clock64();
        const uint32_t idesc = (mode == 4) ? ((1u << 4) | ((uint32_t)(256 >> 3) << 17) | ((uint32_t)(128 >> 4) << 24)) : IDESC;
        for (int it = 0; it < iters; it++) {
#pragma unroll 6
            for (int j = 0; j < per; j++) {
                const uint32_t koff = (uint32_t)((j & 3) * 2) * 2048;
                const uint64_t db = desc(b + koff, 2048, 128), da = desc(a + koff, 2048, 128);
                if (el) asm volatile("{\n.reg .pred p;\nsetp.ne.b32 p, %4, 0;\ntcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n}" ::"r"(tm), "l"(da), "l"(db), "r"(idesc), "r"(1u) : "memory");
            }
            if (it == iters - 1 || mode == 5) {
                if (el) asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(su32(&bar)) : "memory");
                uint32_t ok = 0;
                while (!ok) asm volatile("{\n.reg .pred P1;\nmbarrier.try_wait.parity.shared::cta.b64 P1, [%1], %2;\nselp.u32 %0, 1, 0, P1;\n}" : "=r"(ok) : "r"(su32(&bar)), "r"(phase) : "memory");
                phase ^= 1;
            }
            __syncwarp();
        }
        long long t1 = clock64();
        if (blockIdx.x == 0 && lane == 0) out[0] = t1 - t0;
    } else if (threadIdx.x == 0) {
        const uint32_t a = su32(smem), b = su32(smem + 32768);
        uint32_t phase = 0;
        long long t0 = clock64();
        for (int it = 0; it < iters; it++) {
            for (int j = 0; j < per; j++) {
                const uint32_t koff = (uint32_t)((j & 3) * 2) * 2048; // walk 4 k-steps of a 16 KB operand
                const uint64_t db = desc(b + koff, 2048, 128);
                if (mode == 1) {
                    asm volatile("{\n.reg .pred p;\nsetp.ne.b32 p, %4, 0;\ntcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n}" ::"r"(tm), "r"(tm + 256 + (j & 3) * 8), "l"(db), "r"(IDESC), "r"(1u) : "memory");
                } else {
                    const uint64_t da = desc(a + koff, 2048, 128);
                    asm volatile("{\n.reg .pred p;\nsetp.ne.b32 p, %4, 0;\ntcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n}" ::"r"(tm), "l"(da), "l"(db), "r"(IDESC), "r"(1u) : "memory");
                }
            }
            if (mode == 2 || it == iters - 1) {
                asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(su32(&bar)) : "memory");
                uint32_t ok = 0;
                while (!ok) asm volatile("{\n.reg .pred P1;\nmbarrier.try_wait.parity.shared::cta.b64 P1, [%1], %2;\nselp.u32 %0, 1, 0, P1;\n}" : "=r"(ok) : "r"(su32(&bar)), "r"(phase) : "memory");
                phase ^= 1;
            }
        }
        long long t1 = clock64();
        if (blockIdx.x == 0) out[0] = t1 - t0;
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tm), "r"(512u) : "memory");
}
int main() {
    long long *d, h;
    cudaMalloc(&d, 8);
    cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536);
    for (int grid : {148})
        for (int mode = 0; mode < 6; mode++)
            for (int per : {6, 24}) {
                const int iters = 768 / per;
                k<<<grid, 128, 65536>>>(mode, iters, per, d);
                cudaError_t e = cudaDeviceSynchronize();
                cudaMemcpy(&h, d, 8, cudaMemcpyDeviceToHost);
                printf("grid %3d mode %d per %2d: %s  %.1f cycles per MMA (ideal 64)\n", grid, mode, per, cudaGetErrorString(e), (double)h / 768.0);
            }
    return 0;
}
