// mllm/backends/cuda/CUDABackend.hpp -- the CUDA (B200, sm_100a) backend as it sits INSIDE an mllm checkout.
//
// Compiled against the reference's real headers (mllm/Backend.hpp, Op.hpp, Tensor.hpp, Module.hpp, Layer.hpp): the overlay in
// integration/ is copied next to mllm/backends/cpu, integration/patches/mllm_cuda.patch adds the enum value / initBackend case /
// tensor migration, and unchanged model headers (models/llama/modeling_llama.hpp ...) run on it after `model.to(MLLM_CUDA)`.
//
// Memory model: tensors of this backend keep TensorImpl::location_ == ON_HOST and address CUDA *unified* memory through
// host_ptr_, handed out by CUDAMemoryManager (a size-class pool over cudaMallocManaged blocks that are made device-resident at
// creation).  Every piece of the reference's tensor machinery that is pointer arithmetic -- child views, shape offsets,
// allocFromTemplate's KV-cache aliasing (Tensor.cpp:225-275), aggregated splits -- therefore works unchanged on device memory,
// and host code that still dereferences a tensor (a tokenizer reading logits, CPUKVCache zero-filling its cache at construction)
// stays legal.  All arithmetic runs in the kernels of libmllmcu.so, reached only through the C ABI of include/mllmcu.h.
//
// Ops: compute ops are CUDA classes (CUDAOps.cpp).  Ops that only edit tensor METADATA (F_VIEW, KVCACHE under flash attention,
// where producers write straight into the cache slice and execute() just advances cache_seq_len_) are the reference's own
// classes, constructed on this backend.  Anything else is refused loudly: opCreate throws instead of returning nullptr, so the
// reference's silent CPU fallback (Layer.hpp:214-218) never happens.
#ifndef MLLM_CUDABACKEND_H
#define MLLM_CUDABACKEND_H

#include "Backend.hpp"
#include "MemoryManager.hpp"
#include "Op.hpp"
#include "Types.hpp"
#include <map>
#include <memory>
#include <string>
#include <unordered_map>
#include <vector>

namespace mllm {
class Module;
class Layer;

// Size-class pool over CUDA unified memory (replaces MemoryPoolManager for this backend; mllm/memory/MemoryPoolManager.hpp).  Blocks up
// to a quarter slab are carved from 128 MiB slabs (one cudaMallocManaged / cudaMalloc per slab), so a first forward -- hundreds of new
// tensor sizes -- costs a handful of driver allocations instead of one each; freed blocks go to the size-class free lists.
class CUDAMemoryManager final : public MemoryManager {
public:
    explicit CUDAMemoryManager(int device) : device_(device) {}
    ~CUDAMemoryManager() override;
    void alloc(void **ptr, size_t size, size_t alignment) override;
    void free(void *ptr) override;
    void reserve(size_t bytes); // unified-memory slabs made ahead of use (start-up, not the first request)
    size_t bytes_allocated = 0, pool_hits = 0, pool_misses = 0, driver_calls = 0;
    double driver_ms = 0; // time spent inside driver allocations (and the population of the new range)
    // Weights and the decode plan's buffers are never dereferenced by host code: while this is set, blocks come from plain device
    // memory (cudaMalloc: large pages, no unified-memory bookkeeping on the streaming path) instead of unified memory.
    bool device_only = false;

private:
    int device_;
    std::multimap<size_t, void *> free_list_[2];    // [unified | device-only] size class -> free blocks (stream-ordered reuse: one stream per backend)
    std::unordered_map<void *, std::pair<size_t, int>> live_; // block -> (size class, kind)
    struct Slab { char *base; size_t size, used; };
    std::vector<Slab> slabs_[2];
    std::vector<void *> spare_; // reserved unified slabs not yet in use
    void *raw(size_t bytes, int kind);
    std::vector<void *> owned_; // what the destructor returns to the driver: slabs and the blocks too large for one
};

// What the decode-step fuser reads from an op it folds into a fused launch (pointers stay owned by the op).
struct CUDAFuseInfo {
    const void *weight = nullptr;
    int wtype = 0, in = 0, out = 0;
    const float *bias = nullptr;
    const float *norm_w = nullptr;
    float eps = 0.f;
    int unit_offset = 0;
    const float *sin_t = nullptr, *cos_t = nullptr;
    int rope_type = 0, rot_dim = 0, pos_max = 0;
    float rope_theta = 0.f; // MultimodalRoPE: the decode plan builds the (text-position) table itself
    int mrope = 0;
};
class CUDAOpBase : public Op { // CUDA compute ops: adds the fusion hook
public:
    using Op::Op;
    virtual bool fuseInfo(CUDAFuseInfo &) { return false; }
    virtual void advance(int /*tokens*/) {}
};

class CUDABackend final : public Backend {
public:
    class Creator { // mllm/backends/cpu/CPUBackend.hpp:23-35
    public:
        virtual ~Creator() = default;
        virtual Op *create(OpParam op_param, Backend *bn, string name, int threadCount) const = 0;
    };
    explicit CUDABackend(const BackendConfig &config, int device = 0);
    ~CUDABackend() override;

    Op *opCreate(const OpParam &op_param, string name = "", int threadCount = 4) override;
    TensorFunction *funcCreate(TensorFuncType type) override;
    void registerOps() override;
    void registerVisionOps(); // CUDAVisionOps.cpp
    void registerFuncs() override {}
    std::vector<Tensor> runLayer(Layer *layer, std::vector<Tensor> inputs, int N) override;
    std::vector<Tensor> runOp(Op *op, std::vector<Tensor> input, std::vector<std::string> out_names, bool in_place) override;
    std::vector<Tensor> runForward(Module *module, std::vector<Tensor> inputs, std::vector<std::any> args) override;

    void alloc_device(DeviceMemory &mem, DataType dtype) override;
    void free_device(DeviceMemory &mem) override;
    void copy_from_host(const DeviceMemory &dest, const void *src) override;
    void copy_to_host(void *dest, const DeviceMemory &src) override;
    void convert_fp_data(Tensor *src, Tensor *dest) override;
    bool load_from_file(Tensor *tensor, ParamLoader *loader) override;

    bool addCreator(OpType t, Creator *c);
    void *stream() const { return stream_; }
    int device() const { return device_; }
    CUDAMemoryManager *memory() const { return static_cast<CUDAMemoryManager *>(mem_manager_.get()); }
    void sync();
    void settleKVCache(Op *op); // CUDAOps.cpp: a cache the host has just zero-filled goes back to the device
    void *workspace(size_t bytes); // scratch shared by ops (activation quantisation, split-K partials); grows monotonically
    size_t workspaceBytes() const { return ws_bytes_; }
    // q/k/v and gate/up read one activation tensor back to back: the later Linears reuse the first one's quantised image
    struct QuantTag { const void *x = nullptr; int M = 0, K = 0, ld = 0, kind = 0; };
    bool quantTagMatches(const void *x, int M, int K, int ld, int kind) const {
        return kind != 0 && qtag_.kind == kind && qtag_.x == x && qtag_.M == M && qtag_.K == K && qtag_.ld == ld;
    }
    void setQuantTag(const void *x, int M, int K, int ld, int kind) { qtag_ = QuantTag{x, M, K, ld, kind}; }
    void dropQuantTag() { qtag_ = QuantTag{}; }
    long long kernel_launches = 0;

    // ---- fused single-token step: the op sequence of an S == 1 forward is recorded once, recognised as a LLaMA-family decoder
    // and replayed as the fused kernels of the decode plan (CUDADecodePlan.cpp) without touching the model headers
    struct DecodePlan;
    bool fused_decode = true; // MLLM_CUDA_EAGER_DECODE=1 turns it off
    int last_forward_fused = 0;

private:
    int device_ = 0;
    void *stream_ = nullptr;
    std::map<OpType, Creator *> map_creator_;
    std::map<std::string, Op *> probed_; // ops created by Layer::run's support probe, handed to the create that follows
    void *ws_ = nullptr;
    size_t ws_bytes_ = 0;
    QuantTag qtag_;
    std::unordered_map<Op *, void *> settled_caches_;
    void *stage_ = nullptr; // pinned staging for load_from_file
    size_t stage_bytes_ = 0;
    int load_fd_ = -1;
    std::string load_path_;
    // recorder state
    friend struct DecodePlanBuilder;
    std::vector<Op *> rec_ops_;
    bool recording_ = false;
    std::shared_ptr<DecodePlan> plan_; // shared_ptr: the type is complete only in CUDADecodePlan.cpp
    bool try_decode_plan(Module *module, std::vector<Tensor> &inputs, std::vector<std::any> &args, std::vector<Tensor> &out);
};

[[noreturn]] void cuda_fail(int rc, const char *what, const char *file, int line);
#define MLLMCU_CHECK(call)                                                  \
    do {                                                                    \
        int rc__ = (call);                                                  \
        if (rc__ != 0) ::mllm::cuda_fail(rc__, #call, __FILE__, __LINE__);  \
    } while (0)

} // namespace mllm
#endif // MLLM_CUDABACKEND_H
