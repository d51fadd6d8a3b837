// mllm/backends/cuda/CUDAOpHelpers.hpp -- helpers shared by the op files of the CUDA backend: parameter reading, the element strides of
// a BSHD tensor or child view, and parameter tensors loaded into device memory.
#pragma once
#include "CUDABackend.hpp"
#include "ParamLoader.hpp"
#include "Tensor.hpp"

#include <stdexcept>
#include <string>

namespace mllm {

inline CUDABackend *cb(Backend *b) { return static_cast<CUDABackend *>(b); }
inline float getp(const OpParam &p, const char *k, float dflt) {
    auto it = p.find(k);
    return it == p.end() ? dflt : it->second;
}

// element strides of a BSHD tensor or child view (Tensor::offset, Tensor.hpp:264-311)
struct View {
    long long bs, ss; // elements between consecutive batches / sequence positions; (head, dimension) are contiguous
};
inline View strides(Tensor &t) {
    if (t.ctype() != BSHD) throw std::runtime_error("[mllm-cuda] " + t.name() + ": only BSHD tensors");
    const auto m = t.shapeMaster();
    if (m.size() == 4 && t.shapeOffset().size() == 4) return {(long long)m[2] * m[1] * m[3], (long long)m[1] * m[3]};
    return {(long long)t.sequence() * t.head() * t.dimension(), (long long)t.head() * t.dimension()};
}
inline bool dense(Tensor &t) {
    const View v = strides(t);
    return v.ss == (long long)t.head() * t.dimension() && (t.batch() == 1 || v.bs == (long long)t.sequence() * v.ss);
}
inline void *base(Tensor &t) { // address of element (0,0,0,0) of the view
    const size_t esz = t.dtype() == MLLM_TYPE_F16 ? 2 : 4;
    return (char *)t.rawHostPtr() + (size_t)t.offset(0, 0, 0, 0) * esz;
}
inline float *fbase(Tensor &t) {
    if (t.dtype() != MLLM_TYPE_F32) throw std::runtime_error("[mllm-cuda] " + t.name() + ": fp32 tensor expected");
    return (float *)base(t);
}
inline void load_param(Tensor &t, Backend *bn, const string &name, int rows, int cols, AbstructLoader &loader, DataType dflt) {
    t.setBackend(bn);
    t.setName(name);
    t.reshape(1, 1, rows, cols);
    const DataType dt = loader.getDataType(name);
    t.setDtype(dt != MLLM_TYPE_COUNT ? dt : dflt);
    cb(bn)->memory()->device_only = true; // parameters are read by kernels only
    t.alloc();
    cb(bn)->memory()->device_only = false;
    if (dt != MLLM_TYPE_COUNT) loader.load(&t);
}


} // namespace mllm
