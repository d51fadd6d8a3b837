// mllm/backends/cuda/CUDADecodePlan.cpp -- the fused single-token step behind an UNCHANGED model header.
// (first cut: recogniser not wired yet -- every forward runs op by op through runOp)
#include "CUDABackend.hpp"
#include "Module.hpp"
#include "Tensor.hpp"

namespace mllm {
struct CUDABackend::DecodePlan {};
bool CUDABackend::try_decode_plan(Module *, std::vector<Tensor> &, std::vector<Tensor> &) { return false; }
} // namespace mllm
