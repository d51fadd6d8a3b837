// CUDABackend.cpp -- device storage, the stream-ordered device pool, the .mllm loader's
// upload path and the eager runOp of the CUDA backend.
// Reference interfaces: mllm/Backend.hpp:60-120, mllm/TensorImpl.hpp:105-208,
// mllm/memory/MemoryPoolManager.hpp:15-231 (host pool this replaces on the device),
// mllm/ParamLoader.cpp:88-141,157-286, mllm/backends/cpu/CPUBackend.cpp:314-405 (runOp).
#include "mllm_cuda.hpp"

#include <chrono>
#include <cstring>
#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>

namespace mllm {

map<BackendType, std::unique_ptr<Backend>> Backend::global_backends;
DataType Op::no_load_weights_dtype_ = MLLM_TYPE_F32;

void cuda_fail(int rc, const char *what, const char *file, int line) {
    char buf[512];
    snprintf(buf, sizeof buf, "[mllm-cuda] %s failed with code %d at %s:%d", what, rc, file, line);
    throw std::runtime_error(buf); // OpenCL-backend convention: mllm/backends/opencl/OpenCLBackend.cpp:50-55
}

size_t dtype_block_bytes(DataType t) {
    switch (t) {
    case MLLM_TYPE_F32: case MLLM_TYPE_I32: return 4;
    case MLLM_TYPE_F16: return 2;
    case MLLM_TYPE_Q4_0: case MLLM_TYPE_Q4_0_4_4: return 18; // Q4_0_4_4: 72 B per 4 interleaved blocks (DataType.hpp:104-107)
    case MLLM_TYPE_Q8_0: return 34;
    case MLLM_TYPE_Q4_K: return 144;
    case MLLM_TYPE_Q6_K: return 210;
    default: throw std::runtime_error("[mllm-cuda] unsupported dtype");
    }
}
int dtype_block_elems(DataType t) {
    switch (t) {
    case MLLM_TYPE_Q4_0: case MLLM_TYPE_Q8_0: case MLLM_TYPE_Q4_0_4_4: return 32;
    case MLLM_TYPE_Q4_K: case MLLM_TYPE_Q6_K: return 256;
    default: return 1;
    }
}

// ------------------------------------------------------------------ Tensor
TensorImpl::~TensorImpl() {
    if (owns_ && backend_ && device_memory_.handle) backend_->free_device(device_memory_);
}

Tensor::Tensor(int b, int h, int s, int d, Backend *bn, bool do_alloc) : backend_(bn) {
    reshape(b, h, s, d);
    if (do_alloc) alloc();
}

void Tensor::alloc() {
    if (!backend_) throw std::runtime_error("[mllm-cuda] Tensor::alloc without backend");
    const size_t need = cntSize();
    if (impl_ && impl_->owns_ && impl_->device_memory_.size_in_bytes >= need && !is_view_) return;
    impl_ = std::make_shared<TensorImpl>();
    impl_->backend_ = backend_;
    impl_->owns_ = true;
    impl_->device_memory_.size_in_bytes = need;
    backend_->alloc_device(impl_->device_memory_, dtype_);
    off_ = 0;
    is_view_ = false;
    bstride_ = (long long)s_ * h_ * d_;
    sstride_ = (long long)h_ * d_;
}

void Tensor::shallowCopyFrom(const Tensor &master, int s0, int s) {
    impl_ = master.impl_;
    backend_ = master.backend_;
    dtype_ = master.dtype_;
    b_ = master.b_; h_ = master.h_; d_ = master.d_; s_ = s;
    bstride_ = master.bstride_;
    sstride_ = master.sstride_;
    off_ = master.off_ + (long long)s0 * master.sstride_;
    is_view_ = true;
}

void Tensor::aliasInto(const Tensor &master, int s0) {
    if ((long long)h_ * d_ != (long long)master.h_ * master.d_ || b_ != master.b_) throw std::runtime_error("[mllm-cuda] aliasInto: row size / batch mismatch");
    impl_ = master.impl_;
    dtype_ = master.dtype_;
    bstride_ = master.bstride_;
    sstride_ = master.sstride_;
    off_ = master.off_ + (long long)s0 * master.sstride_;
    is_view_ = true;
}

void *Tensor::rawDevPtr() const {
    if (!impl_) return nullptr;
    return (char *)impl_->device_memory_.handle + (size_t)(off_ / dtype_block_elems(dtype_)) * dtype_block_bytes(dtype_);
}

void Tensor::copyFromHost(const void *src) {
    if (!contiguous()) throw std::runtime_error("[mllm-cuda] copyFromHost on a strided view");
    DeviceMemory m = impl_->device_memory_;
    m.handle = rawDevPtr();
    m.size_in_bytes = cntSize();
    backend_->copy_from_host(m, src);
}
void Tensor::copyToHost(void *dst) const {
    if (!contiguous()) throw std::runtime_error("[mllm-cuda] copyToHost on a strided view");
    DeviceMemory m = impl_->device_memory_;
    m.handle = rawDevPtr();
    m.size_in_bytes = cntSize();
    backend_->copy_to_host(dst, m);
}

ErrorCode Op::setUp(vector<shared_ptr<Tensor>> inputs, vector<shared_ptr<Tensor>> outputs) {
    for (auto &o : outputs) { // mllm/Op.hpp:61-68
        o->setDtype(activation_dtype_);
        o->alloc();
    }
    return MLLM_NO_ERROR;
}

// ------------------------------------------------------------- CUDABackend
CUDABackend::CUDABackend(int device) : device_(device) {
    type_ = MLLM_CUDA;
    if (mllmcu_device_count() <= device) throw std::runtime_error("[mllm-cuda] no CUDA device: the CUDA backend has no CPU fallback");
    MLLMCU_CHECK(mllmcu_set_device(device));
    MLLMCU_CHECK(mllmcu_stream_create(&stream_));
    registerOps();
}

CUDABackend::~CUDABackend() {
    mllmcu_stream_sync(stream_);
    for (auto &kv : free_list_) mllmcu_free(kv.second);
    if (ws_) mllmcu_free(ws_);
    mllmcu_stream_destroy(stream_);
}

// Device pool: 256-byte granularity below 1 MiB, 64 KiB granularity above; freed blocks
// go to a size-keyed free list and are reused by the next request of the same class.
// Everything runs on one stream, so reuse is stream-ordered-safe without events.
static size_t size_class(size_t n) {
    if (n == 0) n = 1;
    const size_t g = n <= (1u << 20) ? 256 : (64u << 10);
    return (n + g - 1) / g * g;
}
void CUDABackend::alloc_device(DeviceMemory &mem, DataType) {
    const size_t cls = size_class(mem.size_in_bytes);
    auto it = free_list_.find(cls);
    if (it != free_list_.end()) {
        mem.handle = it->second;
        free_list_.erase(it);
        pool_hits++;
    } else {
        MLLMCU_CHECK(mllmcu_malloc(&mem.handle, cls));
        bytes_allocated += cls;
        pool_misses++;
    }
    mem.type = MEM_TYPE_GENERIC;
    mem.size_in_bytes = cls;
}
void CUDABackend::free_device(DeviceMemory &mem) {
    if (!mem.handle) return;
    free_list_.emplace(mem.size_in_bytes, mem.handle);
    mem.handle = nullptr;
}
void CUDABackend::copy_from_host(const DeviceMemory &dest, const void *src) {
    dropQuantTag();
    MLLMCU_CHECK(mllmcu_memcpy_h2d(dest.handle, src, dest.size_in_bytes, stream_));
    MLLMCU_CHECK(mllmcu_stream_sync(stream_));
}
void CUDABackend::copy_to_host(void *dest, const DeviceMemory &src) {
    MLLMCU_CHECK(mllmcu_memcpy_d2h(dest, src.handle, src.size_in_bytes, stream_));
    MLLMCU_CHECK(mllmcu_stream_sync(stream_));
}
void CUDABackend::sync() { MLLMCU_CHECK(mllmcu_stream_sync(stream_)); }

void *CUDABackend::workspace(size_t bytes) {
    dropQuantTag();
    return linearWorkspace(bytes);
}
void *CUDABackend::linearWorkspace(size_t bytes) {
    if (bytes > ws_bytes_) {
        dropQuantTag();
        MLLMCU_CHECK(mllmcu_stream_sync(stream_));
        if (ws_) MLLMCU_CHECK(mllmcu_free(ws_));
        ws_bytes_ = (bytes + (8u << 20)) / (1u << 20) * (1u << 20);
        MLLMCU_CHECK(mllmcu_malloc(&ws_, ws_bytes_));
    }
    return ws_;
}

// Weight upload in the file's native block layout: the mmap'ed bytes go straight to the
// tensor's device memory (the OpenCL backend's equivalent maps the buffer and freads into
// it, OpenCLBackend.cpp:928-980).
bool CUDABackend::load_from_file(Tensor *tensor, ParamLoader *loader) {
    size_t bytes = 0;
    const uint8_t *src = loader->data(tensor->name(), &bytes);
    if (!src) return false;
    if (bytes != tensor->cntSize()) {
        char buf[256];
        snprintf(buf, sizeof buf, "[mllm-cuda] %s: file has %zu bytes, tensor expects %zu", tensor->name().c_str(), bytes, tensor->cntSize());
        throw std::runtime_error(buf);
    }
    MLLMCU_CHECK(mllmcu_memcpy_h2d(tensor->rawDevPtr(), src, bytes, stream_));
    MLLMCU_CHECK(mllmcu_stream_sync(stream_));
    return true;
}

bool CUDABackend::addCreator(OpType t, Creator *c) {
    if (map_creator_.count(t)) return false;
    map_creator_[t] = c;
    return true;
}

Op *CUDABackend::opCreate(const OpParam &op_param, string name, int threadCount) {
    auto it_t = op_param.find("type");
    if (it_t == op_param.end()) throw std::runtime_error("[mllm-cuda] opCreate: OpParam has no \"type\"");
    const OpType t = (OpType)(int)it_t->second;
    auto it = map_creator_.find(t);
    if (it == map_creator_.end()) {
        // The reference returns nullptr here and Layer::initOp silently falls back to the CPU op
        // (Layer.hpp:214-218); north_star forbids that on this path, so fail loudly.
        char buf[128];
        snprintf(buf, sizeof buf, "[mllm-cuda] OpType %d (%s) has no CUDA implementation; no CPU fallback", (int)t, name.c_str());
        throw std::runtime_error(buf);
    }
    Op *op = it->second->create(op_param, this, name, threadCount);
    op->type() = t;
    return op;
}

// Eager runOp (CPUBackend.cpp:354-405): output shells -> op->reshape -> allocate (a cache-slice
// view when the output name was registered by a KVCache at trace time) -> op->execute.
// Trace mode (:318-349): reshape + setUp only.
vector<Tensor> CUDABackend::runOp(Op *op, vector<Tensor> inputs, vector<string> out_names, bool in_place) {
    vector<shared_ptr<Tensor>> in, out;
    for (auto &t : inputs) in.push_back(shared_ptr<Tensor>(&t, [](Tensor *) {}));
    vector<Tensor> shells(out_names.size());
    for (size_t i = 0; i < out_names.size(); i++) {
        if (in_place && i < inputs.size()) {
            shells[i] = inputs[i];
        } else {
            shells[i] = Tensor(this);
        }
        shells[i].setName(out_names[i]);
        out.push_back(shared_ptr<Tensor>(&shells[i], [](Tensor *) {}));
    }
    op->reshape(in, out);
    if (doTrace) {
        op->setUp(in, out);
        return shells;
    }
    if (!in_place) {
        for (auto &o : out) {
            auto a = cache_alias.find(o->name());
            if (a != cache_alias.end()) {
                // KVCache template: this output lives in the cache slice [cache_seq_len, +S)
                vector<shared_ptr<Tensor>> one{o};
                a->second->setUp({}, one);
            } else if (!o->allocated()) {
                o->setDtype(op->activation_dtype());
                o->alloc();
            }
        }
    }
    op->execute(in, out);
    if (qtag_.x) { // an op that wrote over the tagged activations invalidates their quantised image
        const char *lo = (const char *)qtag_.x, *hi = lo + (size_t)qtag_.M * qtag_.ld * 4;
        for (auto &o : out) {
            if (!o->allocated()) continue;
            const char *p = (const char *)o->rawDevPtr();
            const char *q = p + (size_t)o->batch() * o->head() * o->sequence() * o->dimension() * 4;
            if (p < hi && lo < q) dropQuantTag();
        }
    }
    return shells;
}

void *CUDABackend::funcCreate(int type) {
    throw std::runtime_error("[mllm-cuda] funcCreate(" + std::to_string(type) + "): tensor functions are Ops on this backend (opCreate with the F_* type)");
}

// .half() / .fp32(): dense tensors of equal shape; fp32 -> fp16 rounds to nearest even like MLLM_FP32_TO_FP16 (F16C), fp16 -> fp32 is exact
void CUDABackend::convert_fp_data(Tensor *src, Tensor *dest) {
    if (!src || !dest || src->count() != dest->count() || !src->contiguous() || !dest->contiguous())
        throw std::runtime_error("[mllm-cuda] convert_fp_data: dense tensors of equal size expected");
    const long long n = (long long)src->count();
    const int cols = src->dimension(), rows = (int)(n / cols);
    if (src->dtype() == MLLM_TYPE_F32 && dest->dtype() == MLLM_TYPE_F16)
        MLLMCU_CHECK(mllmcu_f32_to_f16_rows(src->devPtr<float>(), dest->rawDevPtr(), rows, cols, cols, cols, stream_));
    else if (src->dtype() == MLLM_TYPE_F16 && dest->dtype() == MLLM_TYPE_F32)
        MLLMCU_CHECK(mllmcu_f16_to_f32_rows(src->rawDevPtr(), dest->devPtr<float>(), rows, cols, cols, cols, stream_));
    else
        throw std::runtime_error("[mllm-cuda] convert_fp_data: only fp32 <-> fp16"); // CPUBackend.cpp:274-276
    kernel_launches++;
    dropQuantTag();
}

// The outermost Forward is timed on the host; the clock stops only after the stream has drained.
vector<Tensor> CUDABackend::runForward(const std::function<vector<Tensor>()> &forward, double *ms_out) {
    const auto t0 = std::chrono::steady_clock::now();
    vector<Tensor> out = forward();
    if (ms_out) {
        sync();
        *ms_out = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
    }
    return out;
}

// ------------------------------------------------------------- ParamLoader
ParamLoader::ParamLoader(const string &filename) : path_(filename) {
    if (filename.empty()) return;
    int fd = open(filename.c_str(), O_RDONLY);
    if (fd < 0) { perror(("Error opening file: " + filename).c_str()); throw std::runtime_error("[mllm-cuda] cannot open " + filename); }
    struct stat st;
    fstat(fd, &st);
    size_ = (size_t)st.st_size;
    void *p = mmap(nullptr, size_, PROT_READ, MAP_PRIVATE, fd, 0);
    close(fd);
    if (p == MAP_FAILED) throw std::runtime_error("[mllm-cuda] mmap failed for " + filename);
    base_ = (uint8_t *)p;
    const uint8_t *cur = base_;
    auto rd_i32 = [&]() { int32_t v; memcpy(&v, cur, 4); cur += 4; return v; };
    auto rd_u64 = [&]() { uint64_t v; memcpy(&v, cur, 8); cur += 8; return v; };
    if (size_ < 12 || rd_i32() != 20012) { // _MAGIC_NUMBER
        munmap(base_, size_);
        base_ = nullptr;
        throw std::runtime_error("[mllm-cuda] " + filename + ": magic number error");
    }
    const uint64_t index_size = rd_u64();
    const uint8_t *end = cur + index_size;
    if (end > base_ + size_) throw std::runtime_error("[mllm-cuda] " + filename + ": truncated index");
    while (cur < end) {
        const int32_t len = rd_i32();
        string name((const char *)cur, (size_t)len);
        cur += len;
        const uint64_t length = rd_u64();
        const uint64_t offset = rd_u64();
        const int32_t dt = rd_i32();
        if (offset + length > size_) throw std::runtime_error("[mllm-cuda] " + filename + ": tensor " + name + " out of file bounds");
        offsets_[name] = {offset, length};
        data_type_[name] = dt;
    }
}
ParamLoader::~ParamLoader() {
    if (base_) munmap(base_, size_);
}
DataType ParamLoader::getDataType(const string &name) {
    auto it = data_type_.find(name);
    if (it == data_type_.end()) return MLLM_TYPE_COUNT; // ParamLoader.cpp:317 warns and returns COUNT
    return (DataType)it->second;
}
const uint8_t *ParamLoader::data(const string &name, size_t *bytes) const {
    auto it = offsets_.find(name);
    if (it == offsets_.end() || !base_) return nullptr;
    if (bytes) *bytes = (size_t)it->second.second;
    return base_ + it->second.first;
}
vector<string> ParamLoader::getParamNames() const {
    vector<string> v;
    for (auto &kv : offsets_) v.push_back(kv.first);
    return v;
}
bool ParamLoader::load(Tensor *tensor) {
    if (tensor->backend() && tensor->backend()->load_from_file(tensor, this)) return true;
    return false;
}

} // namespace mllm
