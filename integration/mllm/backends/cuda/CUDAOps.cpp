// mllm/backends/cuda/CUDAOps.cpp -- the CUDA ops of the in-tree backend, written against the reference's real Op / Tensor /
// AbstructLoader interfaces.  Each op keeps the reshape / setUp / load / execute semantics of the CPU op it replaces (cited) and
// reaches the device only through the C ABI of include/mllmcu.h.  Tensors are BSHD; a tensor may be a child view of a KV cache
// (allocFromTemplate), so every kernel call passes the view's base pointer and its batch / sequence strides.
#include "CUDABackend.hpp"

#include "Module.hpp"
#include "ParamLoader.hpp"
#include "Tensor.hpp"
#include "backends/cpu/op/CPUKVCache.hpp"
#include "backends/cpu/op/CPUViewFunc.hpp"
#include "CUDAOpHelpers.hpp"
#include "mllmcu.h"

#include <algorithm>
#include <cmath>
#include <cstring>
#include <stdexcept>

namespace mllm {

// ------------------------------------------------------------------ Linear (op/CPULinear.cpp:22-233, compute/Matmul.cpp:77-279)
class CUDALinear final : public CUDAOpBase {
public:
    CUDALinear(Backend *bn, string name, int in_features, int out_features, bool bias) :
        CUDAOpBase(bn, name), in_(in_features), out_(out_features), bias_(bias) {}
    ErrorCode reshape(vector<shared_ptr<Tensor>> in, vector<shared_ptr<Tensor>> out) override {
        if (in[0]->count() == 0 && in[0]->sequence() != 0) { out[0]->reshape(0, 0, 0, 0); return MLLM_NO_ERROR; }
        if (in[0]->head() != 1 || in[0]->dimension() != in_) throw std::runtime_error("[mllm-cuda] Linear " + name() + ": input must be [B,1,S," + std::to_string(in_) + "]");
        out[0]->reshape(in[0]->batch(), in[0]->head(), in[0]->sequence(), out_);
        return MLLM_NO_ERROR;
    }
    ErrorCode load(AbstructLoader &loader) override { // CPULinear.cpp:47-95
        DataType dt = loader.getDataType(name() + ".weight");
        const DataType eff = dt != MLLM_TYPE_COUNT ? dt : Op::noLoadWeightsDtype();
        if (eff != MLLM_TYPE_F32 && eff != MLLM_TYPE_Q4_0 && eff != MLLM_TYPE_Q8_0 && eff != MLLM_TYPE_Q4_K && eff != MLLM_TYPE_Q6_K)
            throw std::runtime_error("[mllm-cuda] Linear " + name() + ": weight dtype " + std::to_string((int)eff) + " has no CUDA kernel");
        load_param(weight_, backend_, name() + ".weight", out_, in_, loader, eff);
        if (bias_) load_param(biasT_, backend_, name() + ".bias", 1, out_, loader, MLLM_TYPE_F32);
        return Op::load(loader);
    }
    ErrorCode execute(vector<shared_ptr<Tensor>> in, vector<shared_ptr<Tensor>> out) override {
        auto *b = cb(backend_);
        Tensor &x = *in[0], &y = *out[0];
        if (x.count() == 0) return MLLM_NO_ERROR;
        const int f16 = y.dtype() == MLLM_TYPE_F16;
        const float *bias = bias_ ? biasT_.hostPtr<float>() : nullptr;
        if (weight_.dtype() == MLLM_TYPE_F32) { // weights the quantizer keeps F32: fp32 GEMM (vec_dot_fp32's products in another order)
            if (!dense(x)) throw std::runtime_error("[mllm-cuda] Linear " + name() + ": F32 weights need a dense fp32 input");
            const int M = x.batch() * x.sequence();
            const bool direct = !f16 && dense(y);
            float *dst = direct ? fbase(y) : (float *)b->workspace((size_t)M * out_ * 4);
            MLLMCU_CHECK(mllmcu_sgemm(fbase(x), in_, 0, weight_.hostPtr<float>(), in_, 1, 0, bias, dst, out_, 0, M, out_, in_, 1, b->stream()));
            b->kernel_launches++;
            if (!direct) { // the output is a slice of a KV cache (fp16 and / or strided): rows go there through the converting / strided copy
                const View ys = strides(y);
                for (int i = 0; i < y.batch(); i++) {
                    const float *src = dst + (size_t)i * y.sequence() * out_;
                    char *yp = (char *)base(y) + (size_t)i * ys.bs * (f16 ? 2 : 4);
                    if (f16) MLLMCU_CHECK(mllmcu_f32_to_f16_rows(src, yp, y.sequence(), out_, out_, ys.ss, b->stream()));
                    else {
                        const int n[4] = {1, y.sequence(), 1, out_};
                        const long long ss[4] = {0, out_, 0, 1}, ds[4] = {0, ys.ss, 0, 1};
                        MLLMCU_CHECK(mllmcu_copy_strided(src, yp, 4, n, ss, ds, b->stream()));
                    }
                    b->kernel_launches++;
                }
            }
            b->dropQuantTag();
            return MLLM_NO_ERROR;
        }
        const View xs = strides(x), ys = strides(y);
        const bool one = dense(x) && (y.batch() == 1 || ys.bs == (long long)y.sequence() * ys.ss);
        const int nbatch = one ? 1 : x.batch();
        const int M = one ? x.batch() * x.sequence() : x.sequence();
        void *ws = b->workspace(mllmcu_linear_workspace_bytes(M, in_));
        const int kind = nbatch == 1 ? mllmcu_linear_quant_kind((int)weight_.dtype(), M, out_, in_, weight_.rawHostPtr()) : 0;
        for (int i = 0; i < nbatch; i++) {
            const float *xp = fbase(x) + (size_t)i * xs.bs;
            char *yp = (char *)base(y) + (size_t)i * ys.bs * (f16 ? 2 : 4);
            const int reuse = b->quantTagMatches(xp, M, in_, (int)xs.ss, kind) ? 1 : 0;
            MLLMCU_CHECK(mllmcu_linear_ex((int)weight_.dtype(), weight_.rawHostPtr(), xp, bias, yp, M, out_, in_, (int)xs.ss, (int)ys.ss, f16, ws,
                                          b->workspaceBytes(), reuse, b->stream()));
            if (kind) b->setQuantTag(xp, M, in_, (int)xs.ss, kind);
            else b->dropQuantTag();
            b->kernel_launches += 2 - reuse;
        }
        return MLLM_NO_ERROR;
    }
    ErrorCode free(vector<shared_ptr<Tensor>>, vector<shared_ptr<Tensor>>) override { weight_.free(); if (bias_) biasT_.free(); return MLLM_NO_ERROR; }
    bool fuseInfo(CUDAFuseInfo &f) override {
        if (!weight_.allocted()) return false;
        f.weight = weight_.rawHostPtr(); f.wtype = (int)weight_.dtype(); f.in = in_; f.out = out_;
        f.bias = bias_ ? biasT_.hostPtr<float>() : nullptr;
        return true;
    }

private:
    int in_, out_;
    bool bias_;
    Tensor weight_, biasT_;
};

// ----------------------------------------------------------------- RMSNorm (op/CPURMSNorm.cpp:31-135)
class CUDARMSNorm final : public CUDAOpBase {
public:
    CUDARMSNorm(Backend *bn, string name, int norm_size, float eps, bool add_unit_offset) :
        CUDAOpBase(bn, name), n_(norm_size), eps_(eps), unit_(add_unit_offset) {}
    ErrorCode reshape(vector<shared_ptr<Tensor>> in, vector<shared_ptr<Tensor>> out) override {
        out[0]->reshape(in[0]->batch(), in[0]->head(), in[0]->sequence(), in[0]->dimension());
        return MLLM_NO_ERROR;
    }
    ErrorCode load(AbstructLoader &loader) override {
        const DataType dt = loader.getDataType(name() + ".weight");
        if (dt != MLLM_TYPE_COUNT && dt != MLLM_TYPE_F32) throw std::runtime_error("[mllm-cuda] RMSNorm " + name() + ": only F32 weights (the quantizer keeps *norm* F32)");
        load_param(weight_, backend_, name() + ".weight", 1, n_, loader, MLLM_TYPE_F32);
        return Op::load(loader);
    }
    ErrorCode execute(vector<shared_ptr<Tensor>> in, vector<shared_ptr<Tensor>> out) override {
        auto *b = cb(backend_);
        if (!dense(*in[0]) || !dense(*out[0])) throw std::runtime_error("[mllm-cuda] RMSNorm " + name() + ": dense tensors expected");
        const int rows = in[0]->batch() * in[0]->sequence() * in[0]->head();
        MLLMCU_CHECK(mllmcu_rmsnorm(fbase(*in[0]), weight_.hostPtr<float>(), fbase(*out[0]), rows, in[0]->dimension(), eps_, unit_, b->stream()));
        b->kernel_launches++;
        return MLLM_NO_ERROR;
    }
    bool fuseInfo(CUDAFuseInfo &f) override {
        if (!weight_.allocted()) return false;
        f.norm_w = weight_.hostPtr<float>(); f.eps = eps_; f.unit_offset = unit_; f.in = n_;
        return true;
    }

private:
    int n_;
    float eps_;
    bool unit_;
    Tensor weight_;
};

// -------------------------------------------------------------------- RoPE (op/CPURoPE.cpp:22-128,200-515)
// Tables are built on the host as the reference builds them (theta in double -> float, t = s * theta in float, libm sin / cos)
// and uploaded once per (pose type, width, theta, positions); the rotation kernel uses the compiled reference's FMA form.
// Rows are built when a position first needs them, 2048 at a time: the reference fills all max_position_embeddings rows on first use
// (LLaMA-2: 16384 x 64 sinf + cosf, 16 ms of host time inside the first prefill) although a request never passes its KV-cache limit.
struct RopeTables {
    Tensor sin_t, cos_t;
    std::vector<float> theta;
    int rows = 0, ready = 0, dim = 0, pose = 0;
    void ensure(CUDABackend *b, int upto) {
        upto = std::min(rows, (upto + 2047) / 2048 * 2048);
        if (upto <= ready) return;
        const int n = upto - ready;
        std::vector<float> s((size_t)n * dim), c((size_t)n * dim);
        for (int p = ready; p < upto; p++)
            for (int i = 0; i < dim / 2; i++) {
                const float ang = p * theta[i];
                const float sv = sinf(ang), cv = cosf(ang);
                const int d0 = pose == LLAMAROPE ? 2 * i : i;
                const int d1 = pose == LLAMAROPE ? 2 * i + 1 : i + dim / 2;
                const size_t at = (size_t)(p - ready) * dim;
                s[at + d0] = sv; c[at + d0] = cv;
                s[at + d1] = sv; c[at + d1] = cv;
            }
        MLLMCU_CHECK(mllmcu_memcpy_h2d(sin_t.hostPtr<float>() + (size_t)ready * dim, s.data(), s.size() * sizeof(float), b->stream()));
        MLLMCU_CHECK(mllmcu_memcpy_h2d(cos_t.hostPtr<float>() + (size_t)ready * dim, c.data(), c.size() * sizeof(float), b->stream()));
        b->sync(); // the staging vectors go away
        ready = upto;
    }
};
static map<string, shared_ptr<RopeTables>> g_rope_tables;

class CUDARoPE final : public CUDAOpBase {
public:
    CUDARoPE(Backend *bn, string name, const OpParam &p) : CUDAOpBase(bn, name) {
        pose_type_ = (int)getp(p, "pose_type", LLAMAROPE);
        rope_theta_ = getp(p, "rope_theta", 10000.f);
        partial_ = getp(p, "partial_rotary_factor", 1.0f);
        pos_max_ = (int)getp(p, "max_position_embeddings", 16384);
        if (pose_type_ != LLAMAROPE && pose_type_ != HFHUBROPE) throw std::runtime_error("[mllm-cuda] RoPE " + name + ": pose_type " + std::to_string(pose_type_) + " has no CUDA kernel");
        theta_type_ = (int)getp(p, "rope_type", 0); // RoPEThetaType: 0 DEFAULT, 1 LLAMA3 (Types.hpp:180-183)
        if (theta_type_ != 0 && theta_type_ != 1) throw std::runtime_error("[mllm-cuda] RoPE " + name + ": rope_type " + std::to_string(theta_type_) + " has no CUDA path");
        if (theta_type_ == 1) {
            l3_factor_ = p.at("factor"); l3_low_ = p.at("low_freq_factor"); l3_high_ = p.at("high_freq_factor");
            l3_old_ctx_ = p.at("original_max_position_embeddings");
        }
    }
    ErrorCode reshape(vector<shared_ptr<Tensor>> in, vector<shared_ptr<Tensor>> out) override {
        out[0]->reshape(in[0]->batch(), in[0]->head(), in[0]->sequence(), in[0]->dimension());
        ishape_ = (int)(in[0]->dimension() * partial_);
        if (!tab_ || tab_dim_ != ishape_) {
            char key[256];
            snprintf(key, sizeof key, "%p-%d-%d-%.9g-%d-%d-%g-%g-%g-%g", (void *)backend_, pose_type_, ishape_, rope_theta_, pos_max_, theta_type_, l3_factor_, l3_low_, l3_high_, l3_old_ctx_);
            auto it = g_rope_tables.find(key);
            if (it == g_rope_tables.end()) it = g_rope_tables.emplace(key, build()).first;
            tab_ = it->second;
            tab_dim_ = ishape_;
        }
        return MLLM_NO_ERROR;
    }
    ErrorCode execute(vector<shared_ptr<Tensor>> in, vector<shared_ptr<Tensor>> out) override {
        auto *b = cb(backend_);
        Tensor &x = *in[0], &y = *out[0];
        if (h_cnt_ + x.sequence() > pos_max_) throw std::runtime_error("[mllm-cuda] RoPE " + name() + ": position exceeds max_position_embeddings");
        if (!dense(x)) throw std::runtime_error("[mllm-cuda] RoPE " + name() + ": dense input expected");
        tab_->ensure(b, h_cnt_ + x.sequence());
        const View ys = strides(y);
        MLLMCU_CHECK(mllmcu_rope(pose_type_, fbase(x), base(y), x.batch(), x.sequence(), x.head(), x.dimension(), ishape_, h_cnt_,
                                 tab_->sin_t.hostPtr<float>(), tab_->cos_t.hostPtr<float>(), ishape_, y.dtype() == MLLM_TYPE_F16, ys.bs, ys.ss, b->stream()));
        b->kernel_launches++;
        h_cnt_ += x.sequence(); // CPURoPE.cpp:517-530
        return MLLM_NO_ERROR;
    }
    void clearCache() override { h_cnt_ = 0; }
    bool fuseInfo(CUDAFuseInfo &f) override {
        if (!tab_) return false;
        tab_->ensure(cb(backend_), h_cnt_ + 2);
        f.sin_t = tab_->sin_t.hostPtr<float>(); f.cos_t = tab_->cos_t.hostPtr<float>();
        f.rope_type = pose_type_; f.rot_dim = ishape_; f.pos_max = pos_max_;
        return true;
    }
    void advance(int tokens) override { // a fused step was replayed: the next one reads the row of position h_cnt_
        h_cnt_ += tokens;
        tab_->ensure(cb(backend_), h_cnt_ + 1);
    }
    int position() const { return h_cnt_; }

private:
    shared_ptr<RopeTables> build() {
        const int dim = ishape_;
        vector<float> theta(dim / 2);
        for (int i = 0; i < dim / 2; i++) theta[i] = (float)(1.0 / pow((double)rope_theta_, 2.0 * i / dim));
        if (theta_type_ == 1) { // _compute_llama3_theta (op/CPURoPE.cpp:34-72), statement for statement in float
            const float low_freq_wavelen = l3_old_ctx_ / l3_low_, high_freq_wavelen = l3_old_ctx_ / l3_high_;
            for (int i = 0; i < dim / 2; i++) {
                const float wavelen = 2 * M_PI / theta[i];
                if (wavelen > low_freq_wavelen) {
                    theta[i] /= l3_factor_;
                } else if (wavelen >= high_freq_wavelen && wavelen <= low_freq_wavelen) {
                    const float smooth_factor = (l3_old_ctx_ / wavelen - l3_low_) / (l3_high_ - l3_low_);
                    theta[i] = (1 - smooth_factor) * (theta[i] / l3_factor_) + smooth_factor * theta[i];
                }
            }
        }
        auto t = std::make_shared<RopeTables>();
        auto mk = [&](Tensor &dst) {
            dst.setBackend(backend_);
            dst.reshape(1, 1, pos_max_, dim);
            dst.setDtype(MLLM_TYPE_F32);
            cb(backend_)->memory()->device_only = true;
            dst.alloc();
            cb(backend_)->memory()->device_only = false;
        };
        mk(t->sin_t);
        mk(t->cos_t);
        t->theta = theta; t->rows = pos_max_; t->dim = dim; t->pose = pose_type_;
        return t;
    }
    int pose_type_, pos_max_, ishape_ = 0, h_cnt_ = 0, tab_dim_ = -1;
    float rope_theta_, partial_;
    int theta_type_ = 0;
    float l3_factor_ = 8.f, l3_low_ = 1.f, l3_high_ = 4.f, l3_old_ctx_ = 8192.f;
    shared_ptr<RopeTables> tab_;
};

// --------------------------------------------------------------- Embedding (op/CPUEmbedding.cpp:30-143)
class CUDAEmbedding final : public CUDAOpBase {
public:
    CUDAEmbedding(Backend *bn, string name, int hidden, int vocab) : CUDAOpBase(bn, name), hidden_(hidden), vocab_(vocab) {}
    ErrorCode reshape(vector<shared_ptr<Tensor>> in, vector<shared_ptr<Tensor>> out) override {
        out[0]->reshape(in[0]->batch(), 1, in[0]->sequence(), hidden_);
        return MLLM_NO_ERROR;
    }
    ErrorCode load(AbstructLoader &loader) override {
        const DataType dt = loader.getDataType(name() + ".weight");
        const DataType eff = dt != MLLM_TYPE_COUNT ? dt : MLLM_TYPE_Q4_0;
        if (eff != MLLM_TYPE_Q4_0 && eff != MLLM_TYPE_F32) throw std::runtime_error("[mllm-cuda] Embedding " + name() + ": Q4_0 and F32 tables only");
        load_param(weight_, backend_, name() + ".weight", vocab_, hidden_, loader, eff);
        return Op::load(loader);
    }
    ErrorCode execute(vector<shared_ptr<Tensor>> in, vector<shared_ptr<Tensor>> out) override {
        auto *b = cb(backend_);
        const int n = in[0]->batch() * in[0]->sequence();
        if (weight_.dtype() == MLLM_TYPE_F32) MLLMCU_CHECK(mllmcu_embedding_f32(weight_.rawHostPtr(), fbase(*in[0]), fbase(*out[0]), n, hidden_, b->stream()));
        else MLLMCU_CHECK(mllmcu_embedding_q4_0(weight_.rawHostPtr(), fbase(*in[0]), fbase(*out[0]), n, hidden_, b->stream()));
        b->kernel_launches++;
        return MLLM_NO_ERROR;
    }
    bool fuseInfo(CUDAFuseInfo &f) override {
        if (!weight_.allocted()) return false;
        f.weight = weight_.rawHostPtr(); f.wtype = (int)weight_.dtype(); f.in = vocab_; f.out = hidden_;
        return true;
    }

private:
    int hidden_, vocab_;
    Tensor weight_;
};

// ------------------------------------------------------ SiLU, x + y, x * y (op/CPUSiLU.cpp, op/CPUBinaryFunc.hpp)
class CUDAElementwise final : public CUDAOpBase {
public:
    enum Kind { SILU_K, ADD_K, MUL_K };
    CUDAElementwise(Backend *bn, string name, Kind k) : CUDAOpBase(bn, name), kind_(k) {}
    ErrorCode reshape(vector<shared_ptr<Tensor>> in, vector<shared_ptr<Tensor>> out) override {
        int B = in[0]->batch();
        if (in.size() > 1) B = std::max(B, in[1]->batch()); // a batch-1 operand is repeated over the other's batches (CPUBinaryFunc.hpp:220-245)
        out[0]->reshape(B, in[0]->head(), in[0]->sequence(), in[0]->dimension());
        return MLLM_NO_ERROR;
    }
    ErrorCode execute(vector<shared_ptr<Tensor>> in, vector<shared_ptr<Tensor>> out) override {
        auto *b = cb(backend_);
        for (auto &t : in)
            if (!dense(*t)) throw std::runtime_error("[mllm-cuda] " + name() + ": dense operands expected");
        b->dropQuantTag();
        if (kind_ == SILU_K) {
            MLLMCU_CHECK(mllmcu_silu(fbase(*in[0]), fbase(*out[0]), (size_t)in[0]->count(), b->stream()));
            b->kernel_launches++;
            return MLLM_NO_ERROR;
        }
        Tensor &x = *in[0], &y = *in[1], &o = *out[0];
        const size_t per = (size_t)x.head() * x.sequence() * x.dimension();
        if ((size_t)y.head() * y.sequence() * y.dimension() != per || (x.batch() != y.batch() && x.batch() != 1 && y.batch() != 1))
            throw std::runtime_error("[mllm-cuda] " + name() + ": operands must agree off the batch axis (head broadcast has no CUDA path)");
        auto run = [&](const float *a, const float *c, float *d, size_t n) {
            if (kind_ == ADD_K) MLLMCU_CHECK(mllmcu_add(a, c, d, n, b->stream()));
            else MLLMCU_CHECK(mllmcu_mul(a, c, d, n, b->stream()));
            b->kernel_launches++;
        };
        if (x.batch() == y.batch()) run(fbase(x), fbase(y), fbase(o), per * x.batch());
        else
            for (int n = 0; n < o.batch(); n++) // min(n, batch - 1) of the reference: the single batch of the smaller operand
                run(fbase(x) + (x.batch() == 1 ? 0 : n * per), fbase(y) + (y.batch() == 1 ? 0 : n * per), fbase(o) + n * per, per);
        return MLLM_NO_ERROR;
    }

private:
    Kind kind_;
};

// ----------------------------------------------------- F_CLIP (op/CPUClipFunc.hpp:34-140): sequence / batch slices as device copies
class CUDAClip final : public CUDAOpBase {
public:
    CUDAClip(Backend *bn, string name, const OpParam &p) : CUDAOpBase(bn, name) {
        auto rd = [&](const char *k, vector<int> &v) {
            const int n = (int)getp(p, (string(k) + "_size").c_str(), 0);
            for (int i = 0; i < n; i++) v.push_back((int)p.at(string(k) + "_" + std::to_string(i)));
        };
        rd("b", b_); rd("h", h_); rd("s", s_); rd("d", d_);
        if (!h_.empty() || !d_.empty() || !b_.empty()) throw std::runtime_error("[mllm-cuda] clip " + name + ": only sequence slices have a CUDA path");
    }
    ErrorCode reshape(vector<shared_ptr<Tensor>> in, vector<shared_ptr<Tensor>> out) override {
        int S = in[0]->sequence();
        if (s_.size() == 2) S = s_[1] - s_[0];
        else if (s_.size() == 1) S = 1;
        out[0]->reshape(in[0]->batch(), in[0]->head(), S, in[0]->dimension());
        out[0]->setDtype(in[0]->dtype());
        return MLLM_NO_ERROR;
    }
    ErrorCode execute(vector<shared_ptr<Tensor>> in, vector<shared_ptr<Tensor>> out) override {
        auto *b = cb(backend_);
        Tensor &x = *in[0], &y = *out[0];
        int s0 = s_.empty() ? 0 : s_[0];
        if (s0 < 0) s0 += x.sequence();
        if (s0 < 0 || s0 + y.sequence() > x.sequence()) throw std::runtime_error("[mllm-cuda] clip " + name() + ": sequence range outside the tensor");
        const View xs = strides(x), ys = strides(y);
        const int esz = x.dtype() == MLLM_TYPE_F16 ? 2 : 4;
        const int n[4] = {x.batch(), y.sequence(), 1, x.head() * x.dimension()};
        const long long ss[4] = {xs.bs, xs.ss, 0, 1}, ds[4] = {ys.bs, ys.ss, 0, 1};
        MLLMCU_CHECK(mllmcu_copy_strided((const char *)base(x) + (size_t)s0 * xs.ss * esz, base(y), esz, n, ss, ds, b->stream()));
        b->kernel_launches++;
        return MLLM_NO_ERROR;
    }

private:
    vector<int> b_, h_, s_, d_;
};

// ----------------------------------- F_FA2 (op/CPUFlashAttention2Func.hpp:49-125 -> compute/FlashAttention2.hpp:2236-2284)
class CUDAFlashAttention2 final : public CUDAOpBase {
public:
    CUDAFlashAttention2(Backend *bn, string name, bool causal) : CUDAOpBase(bn, name), causal_(causal) {}
    ErrorCode reshape(vector<shared_ptr<Tensor>> in, vector<shared_ptr<Tensor>> out) override {
        out[0]->reshape(in[0]->batch(), in[0]->head(), in[0]->sequence(), in[0]->dimension());
        return MLLM_NO_ERROR;
    }
    ErrorCode execute(vector<shared_ptr<Tensor>> in, vector<shared_ptr<Tensor>> out) override {
        auto *b = cb(backend_);
        Tensor &q = *in[0], &k = *in[1], &v = *in[2], &o = *out[0];
        const int B = q.batch(), Sq = q.sequence(), Sk = k.sequence(), Hq = q.head(), Hkv = k.head(), D = q.dimension();
        if (!dense(q) || !dense(o)) throw std::runtime_error("[mllm-cuda] flash attention: dense q / o expected");
        const View ks = strides(k), vs = strides(v);
        if (ks.ss != (long long)Hkv * D || vs.ss != ks.ss || vs.bs != ks.bs) throw std::runtime_error("[mllm-cuda] flash attention: k / v must share one [B][limit][Hkv][D] layout");
        const size_t wsb = k.dtype() == MLLM_TYPE_F16 ? mllmcu_flash_attn_workspace_bytes(B, Sq, Hq, D) : mllmcu_flash_attn_f32kv_workspace_bytes(B, Sq, Sk, Hq, Hkv, D);
        void *ws = b->workspace(wsb);
        b->dropQuantTag();
        if (k.dtype() == MLLM_TYPE_F16)
            MLLMCU_CHECK(mllmcu_flash_attn_f16kv(fbase(q), base(k), base(v), fbase(o), B, Sq, Sk, Hq, Hkv, D, causal_, ks.bs, ws, b->workspaceBytes(), b->stream()));
        else
            MLLMCU_CHECK(mllmcu_flash_attn_f32kv(fbase(q), fbase(k), fbase(v), fbase(o), B, Sq, Sk, Hq, Hkv, D, causal_, ks.bs, ws, b->workspaceBytes(), b->stream()));
        b->kernel_launches += 2;
        return MLLM_NO_ERROR;
    }

private:
    bool causal_;
};

// ------------------------------------------------------------------ creators
#define CREATOR(NAME, EXPR)                                                                                   \
    class NAME final : public CUDABackend::Creator {                                                          \
    public:                                                                                                   \
        Op *create(OpParam p, Backend *bn, string name, int threadCount) const override { return EXPR; }     \
    }
CREATOR(LinearCreator, new CUDALinear(bn, name, (int)p["in_features"], (int)p["out_features"], (bool)p["bias"]));
CREATOR(RMSNormCreator, new CUDARMSNorm(bn, name, (int)p["norm_size"], p["epsilon"], p.count("add_unit_offset") ? (bool)p["add_unit_offset"] : false));
CREATOR(RoPECreator, new CUDARoPE(bn, name, p));
CREATOR(EmbeddingCreator, new CUDAEmbedding(bn, name, (int)p["hidden_size"], (int)p["vocab_size"]));
CREATOR(SiLUCreator, new CUDAElementwise(bn, name, CUDAElementwise::SILU_K));
CREATOR(AddCreator, new CUDAElementwise(bn, name, CUDAElementwise::ADD_K));
CREATOR(MulCreator, new CUDAElementwise(bn, name, CUDAElementwise::MUL_K));
CREATOR(ClipCreator, new CUDAClip(bn, name, p));
CREATOR(FA2Creator, new CUDAFlashAttention2(bn, name, p.count("causal_mask") ? p["causal_mask"] != 0.f : true));
// metadata-only ops: the reference's own classes, constructed on this backend (they do no arithmetic and, under flash attention,
// touch no tensor data after construction: CPUViewFunc.hpp:31-50, CPUKVCache.cpp:137-151,253-275)
CREATOR(ViewCreator, new CPUviewFunction(bn, name, threadCount, (int)p.at("b"), (int)p.at("h"), (int)p.at("s"), (int)p.at("d")));
// The reference's KVCache zero-fills its cache with a HOST memset in its constructor and in its batch-change reshape (CPUKVCache.cpp:55-66,
// :85-96), which moves the unified-memory pages to the host; they are sent back before a kernel touches them (measured on LLaVA-1.5-7B: 0.6 ms
// per K append and 0.9 ms per V projection of page faults in every block of the first prefill).  Called after construction and after reshape.
void CUDABackend::settleKVCache(Op *op) {
    auto *kv = dynamic_cast<CPUKVCache *>(op);
    void *at = (kv && kv->cache_) ? kv->cache_->rawHostPtr() : nullptr;
    if (at == nullptr) return;
    void *&seen = settled_caches_[op];
    if (seen == at) return;
    MLLMCU_CHECK(mllmcu_mem_prefetch(at, kv->cache_->cntSize(), device_, stream_));
    seen = at;
}
class KVCacheCreator final : public CUDABackend::Creator {
public:
    Op *create(OpParam p, Backend *bn, string name, int threadCount) const override {
        const int n_rep = (int)p["n_rep"];
        const bool fa2 = (bool)p["fa2"];
        if (!fa2 && n_rep > 1) throw std::runtime_error("[mllm-cuda] KVCache " + name + ": eager GQA head replication copies on the host; use attn_implementation flash_attention_2");
        auto *ret = new CPUKVCache(bn, name, (int)p["hidden"], (int)p["head"], n_rep, fa2, (int)p["cache_max"], threadCount);
        ret->setForXnn((bool)p["for_xnn"]);
        cb(bn)->settleKVCache(ret);
        return ret;
    }
};

void CUDABackend::registerOps() {
    addCreator(LINEAR, new LinearCreator());
    addCreator(RMSNORM, new RMSNormCreator());
    addCreator(ROPE, new RoPECreator());
    addCreator(EMBEDDING, new EmbeddingCreator());
    addCreator(SILU, new SiLUCreator());
    addCreator(F_TTADD, new AddCreator());
    addCreator(F_TTMUL, new MulCreator());
    addCreator(F_CLIP, new ClipCreator());
    addCreator(F_FA2, new FA2Creator());
    addCreator(F_VIEW, new ViewCreator());
    addCreator(KVCACHE, new KVCacheCreator());
    registerVisionOps(); // CUDAVisionOps.cpp: the ops the ViT towers and the image splice add
}

} // namespace mllm
