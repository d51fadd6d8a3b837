// p2p.cu -- the exchange step of the optional tensor-parallel Linear (column-parallel q/k/v/gate/up, row-parallel
// o/down): the all-reduce of the row-parallel partial sums, for the decode shape (a few rows of `hidden` floats -- 16 KiB for the 7B),
// as ONE kernel over peer memory instead of a library collective.  NCCL's all-reduce costs ~10-20 us of latency at this size, 64
// times per token; here every rank owns a buffer its peers map through CUDA IPC, and one launch per rank
//   1. tells every peer "my partial for this site is in place" (a release store of the step's epoch into the peer's flag word),
//   2. waits until every peer has said the same (acquire loads of its own, local flag words),
//   3. reads all `world` partials straight out of the peers' memory over NVLink (rank order 0..world-1 on every rank, so all ranks
//      compute bit-identical sums) and adds the residual -- the Linear's "+ residual" epilogue moves here.
// The partial was written by the producing GEMV earlier on the same stream, so it is complete when this kernel starts; the fence
// before the flag store makes it visible system-wide.  Every call site of a step has its own flags and buffer (a "site"), and a site
// is rewritten only one step later -- after this rank passed a later site's wait, which every peer answers only after it finished
// reading this one (stream order) -- so no double buffering is needed as long as a step has two or more sites.  Epochs come from a
// device counter bumped once per step (graph-replayable: nothing per-step is a kernel argument).
// Prefill-sized reductions (hundreds of rows) are bandwidth-bound and go through NCCL (mllm_b200/tp.py).
// The same exchange also runs as the TAIL of the producing GEMV (gemv_tma.cu, mllmcu_gemv_desc.xchg_*): the last CTA of the Linear to
// finish does steps 1-3, so "row-parallel Linear + all-reduce + residual" is one launch; this standalone kernel serves producers that
// are not that GEMV (batches on the tensor-core kernel, tests).
#include "common.cuh"
#include "mllmcu.h"
#include "p2p.cuh"

MLLMCU_API int mllmcu_ipc_get_handle(void *dev_ptr, void *handle64) {
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "ipc handle size");
    CU_TRY(cudaIpcGetMemHandle((cudaIpcMemHandle_t *)handle64, dev_ptr));
    return 0;
}
MLLMCU_API int mllmcu_ipc_open(const void *handle64, void **dev_ptr) {
    cudaIpcMemHandle_t h;
    memcpy(&h, handle64, 64);
    CU_TRY(cudaIpcOpenMemHandle(dev_ptr, h, cudaIpcMemLazyEnablePeerAccess));
    return 0;
}
MLLMCU_API int mllmcu_ipc_close(void *dev_ptr) {
    CU_TRY(cudaIpcCloseMemHandle(dev_ptr));
    return 0;
}

__global__ void k_p2p_bump(unsigned *ctr) { *ctr += 1; }
MLLMCU_API int mllmcu_p2p_bump(unsigned *epoch_ctr, void *stream) {
    k_p2p_bump<<<1, 1, 0, (cudaStream_t)stream>>>(epoch_ctr);
    CU_LAUNCH_CHECK();
    return 0;
}

// one CTA; n % 4 == 0.  peers.buf[r] / peers.flags[r] already point at this SITE's buffer / flag words of rank r.
__global__ void __launch_bounds__(1024) k_p2p_allreduce(P2PPeers peers, int world, int rank, const unsigned *epoch_ctr, const float *__restrict__ residual,
                                                        float *__restrict__ out, int n) {
    // programmatic dependent launch: the producing GEMV must be complete before its partial is announced; the NEXT kernel (a GEMV that
    // first streams its own weights) may start right away and meets this kernel's result at its own dependency wait
    asm volatile("griddepcontrol.wait;" ::: "memory");
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    p2p_exchange_cta(peers, world, rank, *epoch_ctr, residual, out, n);
}

MLLMCU_API int mllmcu_p2p_allreduce(int world, int rank, const void *const *peer_bases, size_t buf_offset, size_t flag_offset, const unsigned *epoch_ctr,
                                    const float *residual, float *out, int n, int pdl, void *stream) {
    if (world < 1 || world > 8 || rank < 0 || rank >= world || n <= 0 || n % 4 || buf_offset % 16 || flag_offset % 4) return -1;
    P2PPeers p;
    for (int r = 0; r < world; r++) {
        p.buf[r] = (const float *)((const char *)peer_bases[r] + buf_offset);
        p.flags[r] = (unsigned *)((char *)peer_bases[r] + flag_offset);
    }
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(1);
    cfg.blockDim = dim3(1024);
    cfg.stream = (cudaStream_t)stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = pdl ? 1 : 0;
    CU_TRY(cudaLaunchKernelEx(&cfg, k_p2p_allreduce, p, world, rank, epoch_ctr, residual, out, n));
    return 0;
}
