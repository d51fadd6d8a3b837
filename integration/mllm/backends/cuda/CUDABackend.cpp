// mllm/backends/cuda/CUDABackend.cpp -- see CUDABackend.hpp.  Host side only: every CUDA call goes through the C ABI of
// include/mllmcu.h (libmllmcu.so); this file contains no <<<>>> and includes no CUDA header.
#include "CUDABackend.hpp"
#include <chrono>

#include "Layer.hpp"
#include "Module.hpp"
#include "ParamLoader.hpp"
#ifdef MLLM_DEBUG_DUMP
#include "CUDADebugDump.hpp"
#endif
#include "Tensor.hpp"
#include "Timing.hpp"
#include "mllmcu.h"

#include <cstdlib>
#include <cstring>
#include <fcntl.h>
#include <iostream>
#include <stdexcept>
#include <unistd.h>

namespace mllm {

[[noreturn]] void cuda_fail(int rc, const char *what, const char *file, int line) {
    throw std::runtime_error(std::string("[mllm-cuda] ") + what + " failed with code " + std::to_string(rc) + " at " + file + ":" + std::to_string(line));
}

// ------------------------------------------------------------------ memory manager
static size_t size_class(size_t bytes) { // 256 B granules up to 64 KiB, then 1/8-octave steps: bounded waste, few classes
    if (bytes <= 65536) return (bytes + 255) & ~(size_t)255;
    size_t p = 65536;
    while (p < bytes) p <<= 1;
    const size_t step = p >> 4;
    return (bytes + step - 1) / step * step;
}

void *CUDAMemoryManager::raw(size_t n, int kind) {
    const auto t0 = std::chrono::steady_clock::now();
    void *q = nullptr;
    if (kind) MLLMCU_CHECK(mllmcu_malloc(&q, n));
    else MLLMCU_CHECK(mllmcu_malloc_managed(&q, n, device_)); // returns with the range resident on the device
    owned_.push_back(q);
    bytes_allocated += n;
    driver_ms += std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
    driver_calls++;
    return q;
}

// Unified-memory slabs made ahead of the first forward (MLLM_CUDA_POOL_MB, default 1024): creating one costs a driver allocation plus
// the population of 128 MiB of device pages, which belongs to start-up, not to the first request.
void CUDAMemoryManager::reserve(size_t bytes) {
    constexpr size_t SLAB = (size_t)128 << 20;
    for (size_t have = 0; have < bytes; have += SLAB) spare_.push_back(raw(SLAB, 0));
}

void CUDAMemoryManager::alloc(void **ptr, size_t size, size_t /*alignment: blocks are 256 B aligned or better*/) {
    if (size == 0) size = 1;
    const size_t cls = size_class(size);
    const int kind = device_only ? 1 : 0;
    auto it = free_list_[kind].find(cls);
    if (it != free_list_[kind].end()) {
        *ptr = it->second;
        free_list_[kind].erase(it);
        live_[*ptr] = {cls, kind};
        pool_hits++;
        return;
    }
    constexpr size_t SLAB = (size_t)128 << 20;
    void *p = nullptr;
    if (cls > SLAB / 4) p = raw(cls, kind);
    else {
        auto &sl = slabs_[kind];
        if (sl.empty() || sl.back().used + cls > sl.back().size) {
            if (kind == 0 && !spare_.empty()) { sl.push_back({(char *)spare_.back(), SLAB, 0}); spare_.pop_back(); }
            else sl.push_back({(char *)raw(SLAB, kind), SLAB, 0});
        }
        p = sl.back().base + sl.back().used; // classes are multiples of 256 bytes: every carved block is 256-byte aligned
        sl.back().used += cls;
    }
    pool_misses++;
    live_[p] = {cls, kind};
    *ptr = p;
}

void CUDAMemoryManager::free(void *ptr) {
    if (!ptr) return;
    auto it = live_.find(ptr);
    if (it == live_.end()) return; // not ours (e.g. an mmap view handed to setHostPtr)
    free_list_[it->second.second].emplace(it->second.first, ptr);
    live_.erase(it);
}

CUDAMemoryManager::~CUDAMemoryManager() {
    for (void *q : owned_) mllmcu_free(q);
}

// ------------------------------------------------------------------ backend
static shared_ptr<MemoryManager> make_mm(int device) { return std::make_shared<CUDAMemoryManager>(device); }

CUDABackend::CUDABackend(const BackendConfig &, int device) : device_(device) {
    mem_manager_ = make_mm(device);
    type_ = BackendType::MLLM_CUDA;
    if (mllmcu_device_count() <= 0) throw std::runtime_error("[mllm-cuda] no CUDA device: the CUDA backend has no CPU fallback");
    MLLMCU_CHECK(mllmcu_set_device(device));
    MLLMCU_CHECK(mllmcu_stream_create(&stream_));
    const char *pool = getenv("MLLM_CUDA_POOL_MB");
    memory()->reserve((size_t)(pool ? std::max(0, atoi(pool)) : 1024) << 20);
    // the scratch shared by the ops starts large enough for a 7B prefill / a ViT tower's attention planes: growing it in the middle of a
    // forward drains the stream and goes through cudaFree + cudaMalloc (MLLM_CUDA_WS_MB, default 128)
    const char *wsmb = getenv("MLLM_CUDA_WS_MB");
    workspace((size_t)(wsmb ? std::max(1, atoi(wsmb)) : 128) << 20);
    const char *e = getenv("MLLM_CUDA_EAGER_DECODE");
    fused_decode = !(e && e[0] == '1');
    registerOps();
}

CUDABackend::~CUDABackend() {
    for (auto &c : map_creator_) delete c.second;
    if (stream_) mllmcu_stream_sync(stream_);
    if (ws_) mllmcu_free(ws_);
    if (stage_) mllmcu_host_free(stage_);
    if (load_fd_ >= 0) close(load_fd_);
    if (stream_) mllmcu_stream_destroy(stream_);
}

void CUDABackend::sync() { MLLMCU_CHECK(mllmcu_stream_sync(stream_)); }

void *CUDABackend::workspace(size_t bytes) {
    if (bytes > ws_bytes_) {
        sync();
        if (ws_) MLLMCU_CHECK(mllmcu_free(ws_));
        ws_ = nullptr;
        const size_t want = std::max(bytes, ws_bytes_ * 2);
        MLLMCU_CHECK(mllmcu_malloc(&ws_, want));
        ws_bytes_ = want;
        dropQuantTag();
    }
    return ws_;
}

bool CUDABackend::addCreator(OpType t, Creator *c) {
    if (map_creator_.count(t)) { delete c; return false; }
    map_creator_[t] = c;
    return true;
}

Op *CUDABackend::opCreate(const OpParam &op_param, string name, int threadCount) {
    const OpType optype = OpType(op_param.find("type")->second);
    auto it = map_creator_.find(optype);
    if (it == map_creator_.end()) // the reference would fall back to a CPU op silently (Layer.hpp:214-218): not here
        throw std::runtime_error("[mllm-cuda] op type " + std::to_string((int)optype) + " (" + name + ") has no CUDA implementation; the CUDA backend has no CPU fallback");
    // Layer::run probes support with one opCreate (check_op_support, Layer.hpp:232-234) and creates the op with a second one; the
    // probe's object would leak -- for a KV cache that is the whole cache.  While a module is switching backends the probe's op is
    // kept and handed to the create that follows.
    Module *m = Module::llm_model_ptr;
    const bool probing = m && m->doChangeBn && !name.empty();
    if (probing) {
        auto pit = probed_.find(name);
        if (pit != probed_.end()) {
            Op *op = pit->second;
            probed_.erase(pit);
            op->type() = optype;
            return op;
        }
    }
    Op *op = it->second->create(op_param, this, name, threadCount);
    op->type() = optype; // tensor functions reach here through Tensor::runFunc, which does not set it; the decode-step recogniser reads it
    if (probing) probed_[name] = op;
    return op;
}

TensorFunction *CUDABackend::funcCreate(TensorFuncType) {
    throw std::runtime_error("[mllm-cuda] legacy TensorFunction table is not used: every tensor function is an Op created through opCreate (F_* types)");
}

// Backend::alloc_device & friends (Backend.hpp:60-71): generic device blocks for callers that hold a DeviceMemory
void CUDABackend::alloc_device(DeviceMemory &mem, DataType) {
    mem.type = MEM_TYPE_GENERIC;
    mem_manager_->alloc(&mem.handle, mem.size_in_bytes, 256);
}
void CUDABackend::free_device(DeviceMemory &mem) {
    mem_manager_->free(mem.handle);
    mem.handle = nullptr;
}
void CUDABackend::copy_from_host(const DeviceMemory &dest, const void *src) {
    dropQuantTag();
    MLLMCU_CHECK(mllmcu_memcpy_h2d(dest.handle, src, dest.size_in_bytes, stream_));
    sync();
}
void CUDABackend::copy_to_host(void *dest, const DeviceMemory &src) {
    MLLMCU_CHECK(mllmcu_memcpy_d2h(dest, src.handle, src.size_in_bytes, stream_));
    sync();
}

void CUDABackend::convert_fp_data(Tensor *src, Tensor *dest) { // Tensor::half() / Tensor::fp32() (Tensor.hpp:1150-1200)
    const int rows = src->batch() * src->head() * src->sequence(), cols = src->dimension();
    if (src->dtype() == MLLM_TYPE_F32 && dest->dtype() == MLLM_TYPE_F16)
        MLLMCU_CHECK(mllmcu_f32_to_f16_rows(src->hostPtr<float>(), dest->rawHostPtr(), rows, cols, cols, cols, stream_));
    else if (src->dtype() == MLLM_TYPE_F16 && dest->dtype() == MLLM_TYPE_F32)
        MLLMCU_CHECK(mllmcu_f16_to_f32_rows(src->rawHostPtr(), dest->hostPtr<float>(), rows, cols, cols, cols, stream_));
    else
        throw std::runtime_error("[mllm-cuda] convert_fp_data: only fp32 <-> fp16");
    kernel_launches++;
}

// ParamLoader::load offers the tensor here first (ParamLoader.cpp:88-91): file bytes -> pinned staging -> device, native blocks.
bool CUDABackend::load_from_file(Tensor *tensor, ParamLoader *loader) {
    if (loader->getDataType(tensor->name()) == MLLM_TYPE_COUNT) return false;
    const ParamMetadata md = loader->getParamMetadata(tensor->name());
    const std::string path = loader->getParamPath();
    if (load_fd_ < 0 || path != load_path_) {
        if (load_fd_ >= 0) close(load_fd_);
        load_fd_ = open(path.c_str(), O_RDONLY);
        if (load_fd_ < 0) throw std::runtime_error("[mllm-cuda] cannot open " + path);
        load_path_ = path;
    }
    const size_t chunk = (size_t)64 << 20;
    if (!stage_) {
        MLLMCU_CHECK(mllmcu_host_alloc(&stage_, chunk));
        stage_bytes_ = chunk;
    }
    const size_t total = std::min((size_t)md.size, tensor->cntSize());
    char *dst = (char *)tensor->rawHostPtr();
    for (size_t off = 0; off < total; off += chunk) {
        const size_t n = std::min(chunk, total - off);
        size_t got = 0;
        while (got < n) {
            const ssize_t r = pread(load_fd_, (char *)stage_ + got, n - got, (off_t)(md.offset + off + got));
            if (r <= 0) throw std::runtime_error("[mllm-cuda] short read of " + tensor->name());
            got += (size_t)r;
        }
        MLLMCU_CHECK(mllmcu_memcpy_h2d(dst + off, stage_, n, stream_));
        sync(); // the staging buffer is reused
    }
    return true;
}

// ------------------------------------------------------------------ runOp / runLayer / runForward
// Clones of the CPU backend's (backends/cpu/CPUBackend.cpp:295-460): trace mode registers tensors and calls setUp only; eager mode
// creates the output shells, reshape, allocFromTemplate (KV-cache child views, aggregated splits), execute.
static void create_output_tensors(std::vector<std::shared_ptr<Tensor>> &out_tensors, const std::vector<std::string> &out_names, Module *module,
                                  map<std::string, std::shared_ptr<Tensor>> &activation_tensors, Backend *backend) {
    for (const auto &out_name : out_names) {
        auto out_tensor = std::make_shared<Tensor>(backend);
        out_tensor->setName(out_name);
        out_tensor->setModule(module);
        auto it = activation_tensors.find(out_name);
        if (it != activation_tensors.end() && out_name.find("-transpose") == std::string::npos && out_tensor->ctype() != it->second->ctype()) {
            out_tensor->chls() = it->second->chls();
            out_tensor->setCtype(it->second->ctype());
        }
        out_tensors.push_back(out_tensor);
    }
}

std::vector<Tensor> CUDABackend::runOp(Op *op, std::vector<Tensor> inputs, std::vector<std::string> out_names, bool in_place) {
    Module *module = inputs.empty() ? Module::llm_model_ptr : inputs[0].module();
    static map<string, shared_ptr<Tensor>> empty_activation_tensors;
    map<string, shared_ptr<Tensor>> &activation_tensors = module ? module->activation_tensors : empty_activation_tensors;
    if (module && module->doTrace) {
        if (module->tracedFlag) {
            vector<Tensor> results = {};
            for (auto &name : out_names) results.push_back(*activation_tensors[name]);
            return results;
        }
        for (auto &input : inputs) {
            if (input.shouldInGraphs() && activation_tensors.find(input.name()) == activation_tensors.end()) {
                activation_tensors[input.name()] = std::make_shared<Tensor>(op->backend());
                activation_tensors[input.name()]->setName(input.name());
                activation_tensors[input.name()]->setModule(module);
            }
        }
        for (const auto &out_name : out_names) {
            if (activation_tensors.find(out_name) == activation_tensors.end()) {
                activation_tensors[out_name] = std::make_shared<Tensor>(op->backend());
                activation_tensors[out_name]->setName(out_name);
                activation_tensors[out_name]->setModule(module);
            }
        }
        vector<shared_ptr<Tensor>> inPtrs;
        for (auto &input : inputs)
            inPtrs.push_back(input.shouldInGraphs() ? activation_tensors[input.name()] : std::shared_ptr<Tensor>(&input, [](Tensor *) {}));
        vector<shared_ptr<Tensor>> outPtrs = {};
        for (auto &name : out_names) outPtrs.push_back(activation_tensors[name]);
        op->setUp(inPtrs, outPtrs);
        vector<Tensor> results = {};
        for (auto &name : out_names) results.push_back(*activation_tensors[name]);
        return results;
    }
    vector<shared_ptr<Tensor>> input_tensors;
    for (auto &input : inputs) input_tensors.push_back(std::shared_ptr<Tensor>(&input, [](Tensor *) {}));
    vector<shared_ptr<Tensor>> out_tensors;
    if (!in_place) {
        create_output_tensors(out_tensors, out_names, module, activation_tensors, op->backend());
    } else {
        for (size_t i = 0; i < input_tensors.size() && i < out_names.size(); ++i) {
            input_tensors[i]->setName(out_names[i]);
            out_tensors.push_back(input_tensors[i]);
        }
    }
    op->reshape(input_tensors, out_tensors);
    if (op->type() == KVCACHE) settleKVCache(op);
    if (!in_place) {
        for (auto &out_tensor : out_tensors) {
            auto act_it = activation_tensors.find(out_tensor->name());
            auto template_it = act_it != activation_tensors.end() ? act_it->second : nullptr;
            out_tensor->allocFromTemplate(template_it);
        }
    }
    if (recording_) rec_ops_.push_back(op);
    op->execute(input_tensors, out_tensors);
#ifdef MLLM_DEBUG_DUMP
    if (getenv("MLLM_DUMP_OPS")) { sync(); mllm_debug_dump(out_tensors); }
#endif
    vector<Tensor> results;
    for (const auto &out_tensor : out_tensors) results.push_back(*out_tensor);
    return results;
}

std::vector<Tensor> CUDABackend::runLayer(Layer *layer, std::vector<Tensor> inputs, int N) {
    vector<string> out_names;
    const int count = (N > 1) ? N : 1;
    for (int i = 0; i < count; ++i)
        out_names.push_back((N > 1) ? "out-" + layer->op_->name() + "-" + std::to_string(i) : "out-" + layer->op_->name());
    return runOp(layer->op_, inputs, out_names, false);
}

std::vector<Tensor> CUDABackend::runForward(Module *module, std::vector<Tensor> inputs, std::vector<std::any> args) {
    if (mllm::Module::llm_model_ptr && (mllm::Module::llm_model_ptr->doLoad || Module::llm_model_ptr->doChangeBn)) {
        return module->Forward(inputs, args);
    }
    uint64_t time_start = 0;
    const bool outer = (inputs[0].ttype() == TensorType::INPUT_TENSOR);
    if (outer) {
        for (auto &input : inputs) {
            input.setModule(module);
            input.setTtype(TensorType::NORMAL_TENSOR);
        }
        mllm::Module::llm_model_ptr = module;
        if (module->prefilling_token_size_ == 0) module->prefilling_token_size_ = inputs[0].sequence() * inputs[0].batch();
        else if (module->decoding_token_size_ == 0) module->decoding_token_size_ = inputs[0].sequence() * inputs[0].batch();
        time_start = mllm_time_us();
    }
    std::vector<Tensor> output;
    last_forward_fused = 0;
    if (!(outer && try_decode_plan(module, inputs, args, output))) output = module->Forward(inputs, args);
    if (outer) {
        sync(); // the clock must not stop before the device has finished (cf. finishQueue, OpenCLBackend.cpp:1087)
        module->inference_times_.push_back((mllm_time_us() - time_start) / 1000.0F);
    }
    return output;
}

} // namespace mllm
