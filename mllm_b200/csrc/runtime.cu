// runtime.cu -- device/memory/stream entry points of the C ABI (include/mllmcu.h).
// These are what CUDABackend::alloc_device / free_device / copy_from_host /
// copy_to_host / load_from_file reduce to (reference interface: mllm/Backend.hpp:60-71,118).
#include "common.cuh"
#include <string.h>

MLLMCU_API int mllmcu_version(void) { return 100; }
MLLMCU_API int mllmcu_device_count(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) return 0;
    return n;
}
MLLMCU_API int mllmcu_set_device(int dev) { CU_TRY(cudaSetDevice(dev)); return 0; }
MLLMCU_API int mllmcu_device_info(int dev, char *name, int *sm_major, int *sm_minor, int *sm_count, size_t *mem_bytes) {
    cudaDeviceProp p;
    CU_TRY(cudaGetDeviceProperties(&p, dev));
    if (name) { strncpy(name, p.name, 255); name[255] = 0; }
    if (sm_major) *sm_major = p.major;
    if (sm_minor) *sm_minor = p.minor;
    if (sm_count) *sm_count = p.multiProcessorCount;
    if (mem_bytes) *mem_bytes = p.totalGlobalMem;
    return 0;
}
MLLMCU_API int mllmcu_malloc(void **dptr, size_t bytes) { CU_TRY(cudaMalloc(dptr, bytes)); return 0; }
MLLMCU_API int mllmcu_free(void *dptr) { CU_TRY(cudaFree(dptr)); return 0; }
MLLMCU_API int mllmcu_malloc_managed(void **ptr, size_t bytes, int device) {
    CU_TRY(cudaMallocManaged(ptr, bytes, cudaMemAttachGlobal));
    // device-resident from the start: kernels never fault on it; a host touch migrates only the touched pages
    cudaMemAdvise(*ptr, bytes, cudaMemAdviseSetPreferredLocation, device);
    cudaMemPrefetchAsync(*ptr, bytes, device, cudaStreamPerThread);
    cudaStreamSynchronize(cudaStreamPerThread); // resident when this returns: a kernel touching the range while the prefetch is still running faults page by page
    (void)cudaGetLastError(); // advice / prefetch are hints: unsupported configurations keep plain managed behaviour
    return 0;
}
MLLMCU_API int mllmcu_mem_prefetch(void *ptr, size_t bytes, int device, void *stream) {
    if (ptr == nullptr || bytes == 0) return 0;
    cudaMemPrefetchAsync(ptr, bytes, device, (cudaStream_t)stream);
    (void)cudaGetLastError(); // a hint: not managed memory / unsupported configuration -> nothing to do
    return 0;
}
MLLMCU_API int mllmcu_memset(void *dptr, int value, size_t bytes, void *stream) {
    CU_TRY(cudaMemsetAsync(dptr, value, bytes, (cudaStream_t)stream)); return 0;
}
MLLMCU_API int mllmcu_memcpy_h2d(void *dst, const void *src, size_t bytes, void *stream) {
    CU_TRY(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyHostToDevice, (cudaStream_t)stream)); return 0;
}
MLLMCU_API int mllmcu_memcpy_d2h(void *dst, const void *src, size_t bytes, void *stream) {
    CU_TRY(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToHost, (cudaStream_t)stream)); return 0;
}
MLLMCU_API int mllmcu_memcpy_d2d(void *dst, const void *src, size_t bytes, void *stream) {
    CU_TRY(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToDevice, (cudaStream_t)stream)); return 0;
}
MLLMCU_API int mllmcu_host_alloc(void **hptr, size_t bytes) { CU_TRY(cudaHostAlloc(hptr, bytes, cudaHostAllocDefault)); return 0; }
MLLMCU_API int mllmcu_host_free(void *hptr) { CU_TRY(cudaFreeHost(hptr)); return 0; }
MLLMCU_API int mllmcu_stream_create(void **stream) {
    cudaStream_t s;
    CU_TRY(cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking));
    *stream = (void *)s;
    return 0;
}
MLLMCU_API int mllmcu_stream_destroy(void *stream) { CU_TRY(cudaStreamDestroy((cudaStream_t)stream)); return 0; }
MLLMCU_API int mllmcu_stream_sync(void *stream) { CU_TRY(cudaStreamSynchronize((cudaStream_t)stream)); return 0; }

// ---- CUDA events on the caller's stream (device-side timing for Module::profiling-style numbers)
MLLMCU_API int mllmcu_event_create(void **ev) {
    cudaEvent_t e;
    CU_TRY(cudaEventCreate(&e));
    *ev = (void *)e;
    return 0;
}
MLLMCU_API int mllmcu_event_destroy(void *ev) { CU_TRY(cudaEventDestroy((cudaEvent_t)ev)); return 0; }
MLLMCU_API int mllmcu_event_record(void *ev, void *stream) { CU_TRY(cudaEventRecord((cudaEvent_t)ev, (cudaStream_t)stream)); return 0; }
MLLMCU_API int mllmcu_event_sync(void *ev) { CU_TRY(cudaEventSynchronize((cudaEvent_t)ev)); return 0; }
MLLMCU_API int mllmcu_event_elapsed_ms(void *start, void *stop, float *ms) {
    CU_TRY(cudaEventElapsedTime(ms, (cudaEvent_t)start, (cudaEvent_t)stop));
    return 0;
}

// ---- CUDA graphs: a decode step is ~200 small launches; the backend captures it once and
// replays it per token (host dispatch, not the kernels, bounds the eager loop otherwise).
MLLMCU_API int mllmcu_graph_begin_capture(void *stream) {
    CU_TRY(cudaStreamBeginCapture((cudaStream_t)stream, cudaStreamCaptureModeThreadLocal));
    return 0;
}
MLLMCU_API int mllmcu_graph_end_capture(void *stream, void **graph_exec) {
    cudaGraph_t g = nullptr;
    CU_TRY(cudaStreamEndCapture((cudaStream_t)stream, &g));
    cudaGraphExec_t e = nullptr;
    cudaError_t rc = cudaGraphInstantiate(&e, g, 0);
    cudaGraphDestroy(g);
    CU_TRY(rc);
    *graph_exec = (void *)e;
    return 0;
}
// abandon a capture after an error inside it (leaves the stream usable)
MLLMCU_API int mllmcu_graph_abort_capture(void *stream) {
    cudaGraph_t g = nullptr;
    cudaStreamEndCapture((cudaStream_t)stream, &g);
    if (g) cudaGraphDestroy(g);
    cudaGetLastError();
    return 0;
}
MLLMCU_API int mllmcu_graph_launch(void *graph_exec, void *stream) {
    CU_TRY(cudaGraphLaunch((cudaGraphExec_t)graph_exec, (cudaStream_t)stream));
    return 0;
}
MLLMCU_API int mllmcu_graph_destroy(void *graph_exec) {
    if (graph_exec) CU_TRY(cudaGraphExecDestroy((cudaGraphExec_t)graph_exec));
    return 0;
}
