// p2p.cuh -- device side of the peer-memory exchange of the tensor-parallel row-parallel Linear (see p2p.cu), shared by the standalone
// exchange kernel (k_p2p_allreduce) and by the decode GEMV, whose last CTA to finish runs it as the kernel's tail (gemv_tma.cu).
#pragma once
#include <cuda_runtime.h>

struct P2PPeers { const float *buf[8]; unsigned *flags[8]; };

static __device__ __forceinline__ void p2p_st_release_sys(unsigned *p, unsigned v) { asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory"); }
static __device__ __forceinline__ unsigned p2p_ld_acquire_sys(const unsigned *p) {
    unsigned v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
static __device__ __forceinline__ float4 p2p_ld_peer4(const float *p) { // never through L1: the line may hold the previous step's partial
    float4 v;
    asm volatile("ld.relaxed.sys.global.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p) : "memory");
    return v;
}

// Called by ALL threads of one CTA (>= `world` threads), after this rank's partial is complete in peers.buf[rank]:
//   announce it to every peer, wait for every peer's announcement, out[i] = residual[i] + sum_r peers.buf[r][i] in rank order.
static __device__ __forceinline__ void p2p_exchange_cta(const P2PPeers &peers, int world, int rank, unsigned epoch, const float *__restrict__ residual,
                                                        float *__restrict__ out, int n) {
    const int tid = threadIdx.x;
    if (tid < world) {
        __threadfence_system(); // the partial (written by this grid or by an earlier kernel of the stream) before the flag
        p2p_st_release_sys(peers.flags[tid] + rank, epoch);
        const unsigned *mine = peers.flags[rank] + tid;
        unsigned long long t0 = 0;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
        while (p2p_ld_acquire_sys(mine) != epoch) {
            unsigned long long t1;
            asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1));
            if (t1 - t0 > 10000000000ull) __trap(); // 10 s: a peer died or the ranks disagree on the step -- fail, do not hang the box
        }
    }
    __syncthreads();
    for (int i = tid * 4; i < n; i += blockDim.x * 4) {
        float4 acc = residual ? *(const float4 *)(residual + i) : make_float4(0.f, 0.f, 0.f, 0.f);
        float4 v[8];
#pragma unroll
        for (int r = 0; r < 8; r++)
            if (r < world) v[r] = p2p_ld_peer4(peers.buf[r] + i);
#pragma unroll
        for (int r = 0; r < 8; r++)
            if (r < world) { acc.x += v[r].x; acc.y += v[r].y; acc.z += v[r].z; acc.w += v[r].w; }
        *(float4 *)(out + i) = acc;
    }
}
