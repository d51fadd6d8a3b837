// mllm_cuda.hpp -- host side of the CUDA backend: the classes a maintainer would
// place under mllm/backends/cuda/, written against the SAME plugin surface the
// reference defines (mllm/Backend.hpp:32-130, mllm/Op.hpp:20-148,
// mllm/OpDefined.hpp:10-134, mllm/Types.hpp:21-112, mllm/TensorImpl.hpp:23-208).
//
// Inside the mllm tree these ops derive from mllm::Op / mllm::Backend directly
// (INTEGRATION.md shows the three-line edits).  This repository has no mllm
// checkout to compile against on the GPU box, so the header restates the part
// of that surface the hot path touches -- same names, same argument meaning,
// same error behaviour -- and the ops below are written against it.  All
// device work goes through the C ABI in include/mllmcu.h; this layer contains
// no kernels and no CPU compute path.
#pragma once
#include <cstdint>
#include <cstdio>
#include <functional>
#include <map>
#include <memory>
#include <stdexcept>
#include <string>
#include <unordered_map>
#include <vector>

#include "mllmcu.h"

namespace mllm {
using std::map;
using std::shared_ptr;
using std::string;
using std::vector;

// ---- mllm/Types.hpp:34-39 (+ the appended MLLM_CUDA value)
enum BackendType { MLLM_CPU = 0, MLLM_OPENCL, MLLM_QNN, MLLM_XNNPACK, MLLM_CUDA };
// ---- mllm/Types.hpp:54-61
enum ErrorCode { MLLM_NO_ERROR = 0, OUT_OF_MEMORY = 1, NOT_SUPPORT = 2, COMPUTE_SIZE_ERROR = 3, NO_EXECUTION = 4, INVALID_VALUE = 5 };
// ---- mllm/Types.hpp:63-96 (values used on this path)
enum DataType { MLLM_TYPE_F32 = 0, MLLM_TYPE_F16 = 1, MLLM_TYPE_Q4_0 = 2, MLLM_TYPE_Q8_0 = 8, MLLM_TYPE_Q4_K = 12, MLLM_TYPE_Q6_K = 14, MLLM_TYPE_Q8_K = 15, MLLM_TYPE_I32 = 18, MLLM_TYPE_Q4_0_4_4 = 19, MLLM_TYPE_COUNT = 34 };
enum ChlType { BSHD = 0 };
enum RoPEType { NONE = 0, LLAMAROPE = 2, PERSIMMONROPE = 3, HFHUBROPE = 4, MLAROPE = 5, NTKROPE = 6 }; // mllm/Types.hpp:171-178 (only 2 and 4 have kernels; the rest are refused by CUDARoPE)
// ---- mllm/OpDefined.hpp:10-134 (numeric values must match: OpParam["type"] carries them)
enum OpType {
    INVALID_VALUE_OP = 0, PARAMETER = 1, ADD = 2, SOFTMAX = 3, SILU = 4, MATMUL = 6, ROPE = 8, RMSNORM = 11, LINEAR = 14,
    EMBEDDING = 17, MUL = 18, VIEW = 19, KVCACHE = 20, OP_GELU = 24, QUICKGLUE = 25, LAYERNORM = 26, CONVOLUTION2D = 29, CONVOLUTION3D = 30,
    MULTIMODALROPE = 35, F_DIV = 81, F_TTADD = 83, F_TTMUL = 85, F_MM = 87, F_CAT = 90, F_VIEW = 91, F_TRANPOSE = 92, F_CLIP = 94, F_SPLIT = 100,
    F_APPLY_VISIOROPE = 109, F_FA2 = 110,
};
typedef map<string, float> OpParam; // mllm/Types.hpp:21

size_t dtype_block_bytes(DataType t);
int dtype_block_elems(DataType t);

class Backend;
class Op;
class ParamLoader;

// ---- mllm/TensorImpl.hpp:23-45
enum DeviceMemType { MEM_TYPE_GENERIC = 0 };
struct DeviceMemory {
    void *handle = nullptr;
    DeviceMemType type = MEM_TYPE_GENERIC;
    size_t size_in_bytes = 0;
};

// Storage shared by tensor handles (mllm/TensorImpl.hpp:47-144).
struct TensorImpl {
    Backend *backend_ = nullptr;
    DeviceMemory device_memory_;
    bool owns_ = false;
    ~TensorImpl();
};

// 4-D BSHD tensor handle (mllm/Tensor.hpp).  A child view shares its master's
// storage and addresses it with (offset, batch stride, sequence stride), the
// device form of shape_offset_/shape_master_ (Tensor.hpp:264-311).
class Tensor {
public:
    Tensor() = default;
    explicit Tensor(Backend *bn) : backend_(bn) {}
    Tensor(int b, int h, int s, int d, Backend *bn, bool do_alloc = true);

    void reshape(int b, int h, int s, int d) { b_ = b; h_ = h; s_ = s; d_ = d; bstride_ = (long long)s * h * d; sstride_ = (long long)h * d; }
    void alloc();
    void free() { impl_.reset(); off_ = 0; }
    bool allocated() const { return impl_ && impl_->device_memory_.handle; }
    // view of `master` rows [s0, s0+s) (all batches, all heads): mllm's shallowCopyFrom
    void shallowCopyFrom(const Tensor &master, int s0, int s);
    // keep this tensor's logical shape but place it in `master`'s storage at row s0 (H*D must match):
    // what allocFromTemplate does for a producer whose output is a KV-cache child (Tensor.cpp:225-275)
    void aliasInto(const Tensor &master, int s0);
    void reshapeKeepStrides(int b, int h, int s, int d) { b_ = b; h_ = h; s_ = s; d_ = d; }
    // name of the tensor a metadata-only view chain started from (used to register cache aliases)
    const string &origin() const { return origin_.empty() ? name_ : origin_; }
    void setOrigin(const string &o) { origin_ = o; }

    int batch() const { return b_; }
    int head() const { return h_; }
    int sequence() const { return s_; }
    int dimension() const { return d_; }
    size_t count() const { return (size_t)b_ * h_ * s_ * d_; }
    size_t cntSize() const { return count() / dtype_block_elems(dtype_) * dtype_block_bytes(dtype_); }
    DataType dtype() const { return dtype_; }
    void setDtype(DataType t) { dtype_ = t; }
    ChlType ctype() const { return BSHD; }
    const string &name() const { return name_; }
    void setName(const string &n) { name_ = n; }
    Backend *backend() const { return backend_; }
    void setBackend(Backend *b) { backend_ = b; }
    long long batchStride() const { return bstride_; } // elements
    long long seqStride() const { return sstride_; }   // elements
    bool contiguous() const { return sstride_ == (long long)h_ * d_ && (b_ == 1 || bstride_ == (long long)s_ * h_ * d_); }
    bool isView() const { return is_view_; }

    void *rawDevPtr() const;
    template <typename T> T *devPtr() const { return (T *)rawDevPtr(); }
    // synchronous host transfers (Backend::copy_from_host / copy_to_host)
    void copyFromHost(const void *src);
    void copyToHost(void *dst) const;

    DeviceMemory &device_memory() { return impl_->device_memory_; }
    shared_ptr<TensorImpl> impl_;

private:
    Backend *backend_ = nullptr;
    string name_, origin_;
    DataType dtype_ = MLLM_TYPE_F32;
    int b_ = 0, h_ = 0, s_ = 0, d_ = 0;
    long long bstride_ = 0, sstride_ = 0, off_ = 0; // elements of dtype_
    bool is_view_ = false;
};

// What the backend's decode-step fuser reads from an op it folds into a fused launch (device pointers
// stay owned by the op).  No reference counterpart: the CPU backend executes ops one by one.
struct FuseInfo {
    const void *weight = nullptr; // Linear / Embedding: quantised rows in the .mllm layout
    int wtype = 0, in = 0, out = 0;
    const float *bias = nullptr;
    const float *norm_w = nullptr; // RMSNorm
    float eps = 0.f;
    int unit_offset = 0;
    const float *sin_t = nullptr, *cos_t = nullptr; // RoPE tables [pos_max][rot_dim]
    int rope_type = 0, rot_dim = 0, pos = 0, pos_max = 0;
    void *cache = nullptr; // KVCache fp16 [B][limit][H][D]
    long long cache_bstride = 0;
    int cache_limit = 0, cache_len = 0, cache_batch = 0, cache_heads = 0, cache_dim = 0;
};

// ---- mllm/Op.hpp:20-148
class Op {
public:
    Op(Backend *bn, string name = "") : backend_(bn), name_(std::move(name)) {}
    virtual ~Op() = default;
    virtual ErrorCode reshape(vector<shared_ptr<Tensor>> inputs, vector<shared_ptr<Tensor>> outputs) { return MLLM_NO_ERROR; }
    virtual ErrorCode setUp(vector<shared_ptr<Tensor>> inputs, vector<shared_ptr<Tensor>> outputs);
    virtual ErrorCode load(ParamLoader &loader) { return MLLM_NO_ERROR; }
    virtual ErrorCode execute(vector<shared_ptr<Tensor>> inputs, vector<shared_ptr<Tensor>> outputs) { return MLLM_NO_ERROR; }
    virtual ErrorCode free(vector<shared_ptr<Tensor>> inputs, vector<shared_ptr<Tensor>> outputs) { return MLLM_NO_ERROR; }
    virtual int getCacheSeqLen() { return -1; }
    virtual void clearCache() {}
    virtual bool fuseInfo(FuseInfo &) { return false; } // decode-step fusion hook (CUDA backend only)
    virtual void advance(int tokens) {}                 // a fused launch consumed `tokens` positions on this op's behalf
    Backend *backend() const { return backend_; }
    const string &name() const { return name_; }
    OpType &type() { return type_; }
    DataType activation_dtype() const { return activation_dtype_; }
    static DataType &noLoadWeightsDtype() { return no_load_weights_dtype_; }

protected:
    Backend *backend_;
    string name_;
    DataType activation_dtype_ = MLLM_TYPE_F32;
    OpType type_ = INVALID_VALUE_OP;
    static DataType no_load_weights_dtype_;
};

// ---- mllm/Backend.hpp:32-130
class Backend {
public:
    virtual ~Backend() = default;
    virtual void alloc_device(DeviceMemory &mem, DataType dtype) = 0;
    virtual void free_device(DeviceMemory &mem) = 0;
    virtual void copy_from_host(const DeviceMemory &dest, const void *src) = 0;
    virtual void copy_to_host(void *dest, const DeviceMemory &src) = 0;
    virtual Op *opCreate(const OpParam &op_param, string name = "", int threadCount = 4) = 0;
    virtual vector<Tensor> runOp(Op *op, vector<Tensor> inputs, vector<string> out_names, bool in_place) = 0;
    virtual void registerOps() = 0;
    virtual void registerFuncs() {}                              // legacy TensorFunction table (mllm/Backend.hpp:58): nothing to register
    virtual void *funcCreate(int /*TensorFuncType*/) { return nullptr; }
    // .half() / .fp32() between tensors of this backend (mllm/Backend.hpp:73)
    virtual void convert_fp_data(Tensor *src, Tensor *dest) = 0;
    // Layer::run's hand-off (CPUBackend.cpp:407-417): names the N outputs "out-<op>[-i]" and calls runOp
    virtual vector<Tensor> runLayer(class Layer *layer, vector<Tensor> inputs, int N) = 0;
    // Module::operator() hand-off (CPUBackend.cpp:419-460): runs the module's Forward; the outermost call is timed and must not
    // stop the clock before the device has finished (cf. finishQueue, OpenCLBackend.cpp:1087)
    virtual vector<Tensor> runForward(const std::function<vector<Tensor>()> &forward, double *ms_out) = 0;
    virtual bool load_from_file(Tensor *tensor, ParamLoader *loader) { return false; }
    BackendType type() const { return type_; }
    static map<BackendType, std::unique_ptr<Backend>> global_backends; // mllm/Backend.hpp:125

protected:
    BackendType type_ = MLLM_CPU;
};

// ---- .mllm weight container (mllm/ParamLoader.{hpp,cpp}: magic 20012, u64 index length,
// entries {i32 name_len, name, u64 size, u64 offset, i32 dtype}, raw tensor bytes).
class ParamLoader {
public:
    explicit ParamLoader(const string &filename);
    ~ParamLoader();
    bool load(Tensor *tensor);               // offers the tensor to backend->load_from_file first (ParamLoader.cpp:88-141)
    DataType getDataType(const string &name); // MLLM_TYPE_COUNT when absent
    bool has(const string &name) const { return offsets_.count(name) != 0; }
    const uint8_t *data(const string &name, size_t *bytes) const;
    vector<string> getParamNames() const;
    bool ok() const { return base_ != nullptr; }

private:
    string path_;
    uint8_t *base_ = nullptr;
    size_t size_ = 0;
    map<string, std::pair<uint64_t, uint64_t>> offsets_;
    map<string, int> data_type_;
};

// ---- the new backend: class CUDABackend : public Backend (template: mllm/backends/opencl/OpenCLBackend.hpp:26,
// runOp semantics: mllm/backends/cpu/CPUBackend.cpp:314-405)
class CUDABackend : public Backend {
public:
    class Creator { // mllm/backends/cpu/CPUBackend.hpp:23-35
    public:
        virtual ~Creator() = default;
        virtual Op *create(OpParam op_param, Backend *bn, string name, int threadCount) const = 0;
    };
    explicit CUDABackend(int device = 0);
    ~CUDABackend() override;

    void alloc_device(DeviceMemory &mem, DataType dtype) override;
    void free_device(DeviceMemory &mem) override;
    void copy_from_host(const DeviceMemory &dest, const void *src) override;
    void copy_to_host(void *dest, const DeviceMemory &src) override;
    bool load_from_file(Tensor *tensor, ParamLoader *loader) override;
    Op *opCreate(const OpParam &op_param, string name = "", int threadCount = 4) override;
    vector<Tensor> runOp(Op *op, vector<Tensor> inputs, vector<string> out_names, bool in_place) override;
    void registerOps() override;
    void *funcCreate(int type) override; // throws, like OpenCLBackend::funcCreate (OpenCLBackend.cpp:805): every function is an Op here
    void convert_fp_data(Tensor *src, Tensor *dest) override;
    vector<Tensor> runLayer(class Layer *layer, vector<Tensor> inputs, int N) override;
    vector<Tensor> runForward(const std::function<vector<Tensor>()> &forward, double *ms_out) override;
    bool addCreator(OpType t, Creator *c);

    void *stream() const { return stream_; }
    void sync();
    // scratch shared by ops (activation quantisation, split-K partials); grows monotonically
    void *workspace(size_t bytes);
    size_t workspaceBytes() const { return ws_bytes_; }
    // Several Linears reading one activation tensor back to back (q/k/v, gate/up) share its quantised image in the
    // workspace (mllmcu_linear_ex flag 1).  The tag names what the workspace holds; it is dropped by any other workspace
    // user, by a workspace reallocation, by an H2D copy and by any op whose output overlaps the tagged activations.
    struct QuantTag { const void *x = nullptr; int M = 0, K = 0, ld = 0, kind = 0; };
    void *linearWorkspace(size_t bytes);
    bool quantTagMatches(const void *x, int M, int K, int ld, int kind) const {
        return kind != 0 && qtag_.kind == kind && qtag_.x == x && qtag_.M == M && qtag_.K == K && qtag_.ld == ld;
    }
    void setQuantTag(const void *x, int M, int K, int ld, int kind) { qtag_ = QuantTag{x, M, K, ld, kind}; }
    void dropQuantTag() { qtag_ = QuantTag{}; }

    // trace mode (Module::doTrace): runOp only calls reshape + setUp (CPUBackend.cpp:318-349)
    bool doTrace = false;
    // output-name -> producer of a cache slice view (the allocFromTemplate aliasing,
    // Tensor.cpp:225-275 / CPUKVCache.cpp:253-275): filled by CUDAKVCache::setUp
    std::unordered_map<string, Op *> cache_alias;
    // statistics
    size_t bytes_allocated = 0, pool_hits = 0, pool_misses = 0;
    long long kernel_launches = 0; // kernels issued through the C ABI (counted by ops)

private:
    int device_;
    void *stream_ = nullptr;
    map<OpType, Creator *> map_creator_;
    std::multimap<size_t, void *> free_list_; // stream-ordered arena: size class -> free blocks
    void *ws_ = nullptr;
    size_t ws_bytes_ = 0;
    QuantTag qtag_;
};

[[noreturn]] void cuda_fail(int rc, const char *what, const char *file, int line);
#define MLLMCU_CHECK(call)                                              \
    do {                                                                \
        int rc__ = (call);                                              \
        if (rc__ != 0) ::mllm::cuda_fail(rc__, #call, __FILE__, __LINE__); \
    } while (0)

} // namespace mllm
