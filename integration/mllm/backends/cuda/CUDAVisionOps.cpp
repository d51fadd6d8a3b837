// mllm/backends/cuda/CUDAVisionOps.cpp -- the ops the vision towers and the image splice add to the LLaMA-family set of CUDAOps.cpp, so
// that models/llava/modeling_llava.hpp and models/qwen2_vl/modeling_qwen2_vl.hpp run UNCHANGED on this backend:
//   LayerNorm, GELU, QuickGELU, Parameter, Convolution2D, Convolution3D, VisionRoPE, MultimodalRoPE and the tensor functions
//   transpose, flatten, cat, split, range, where, index_put, zero_like, apply_rotary_pos_emb_vision.
// Every tensor of this backend is dense BSHD (Tensor::transpose is never in place off the CPU backend, Tensor.cpp:523-527), so the
// reference's metadata transposes become one strided copy and flatten / view stay metadata.  Each op cites the CPU op whose
// reshape / load / execute semantics it keeps; the device is reached only through the C ABI (include/mllmcu.h).
#include "CUDAOpHelpers.hpp"

#include "Module.hpp"
#include "mllmcu.h"

#include <cmath>
#include <cstring>
#include <map>
#include <stdexcept>
#include <utility>

namespace mllm {

static void need_dense_f32(Tensor &t, const string &who) {
    if (t.dtype() != MLLM_TYPE_F32 || !dense(t)) throw std::runtime_error("[mllm-cuda] " + who + ": dense fp32 tensor expected (" + t.name() + ")");
}
static void same_shape(vector<shared_ptr<Tensor>> &in, vector<shared_ptr<Tensor>> &out) {
    out[0]->reshape(in[0]->batch(), in[0]->head(), in[0]->sequence(), in[0]->dimension());
}
// element strides of a dense BSHD tensor indexed by Chl (BATCH 0, SEQUENCE 1, HEAD 2, DIMENSION 3)
static void chl_geometry(Tensor &t, int n[4], long long st[4]) {
    const View v = strides(t);
    n[BATCH] = t.batch(); n[SEQUENCE] = t.sequence(); n[HEAD] = t.head(); n[DIMENSION] = t.dimension();
    st[BATCH] = v.bs; st[SEQUENCE] = v.ss; st[HEAD] = t.dimension(); st[DIMENSION] = 1;
}

// ------------------------------------------------------------------ LayerNorm (op/CPULayerNorm.cpp:49-89)
class CUDALayerNorm final : public CUDAOpBase {
public:
    CUDALayerNorm(Backend *bn, string name, int norm_size, bool bias, float eps) : CUDAOpBase(bn, name), n_(norm_size), bias_(bias), eps_(eps) {}
    ErrorCode reshape(vector<shared_ptr<Tensor>> in, vector<shared_ptr<Tensor>> out) override { same_shape(in, out); return MLLM_NO_ERROR; }
    ErrorCode load(AbstructLoader &loader) override {
        load_param(weight_, backend_, name() + ".weight", 1, n_, loader, MLLM_TYPE_F32);
        if (bias_) load_param(biasT_, backend_, name() + ".bias", 1, n_, loader, MLLM_TYPE_F32);
        if (weight_.dtype() != MLLM_TYPE_F32 || (bias_ && biasT_.dtype() != MLLM_TYPE_F32)) throw std::runtime_error("[mllm-cuda] LayerNorm " + name() + ": F32 parameters only");
        return Op::load(loader);
    }
    ErrorCode execute(vector<shared_ptr<Tensor>> in, vector<shared_ptr<Tensor>> out) override {
        auto *b = cb(backend_);
        need_dense_f32(*in[0], "LayerNorm");
        need_dense_f32(*out[0], "LayerNorm");
        if (in[0]->dimension() != n_) throw std::runtime_error("[mllm-cuda] LayerNorm " + name() + ": last dimension differs from norm_size");
        const int rows = in[0]->batch() * in[0]->sequence() * in[0]->head();
        b->dropQuantTag();
        MLLMCU_CHECK(mllmcu_layernorm(fbase(*in[0]), weight_.hostPtr<float>(), bias_ ? biasT_.hostPtr<float>() : nullptr, fbase(*out[0]), rows, n_, eps_, b->stream()));
        b->kernel_launches++;
        return MLLM_NO_ERROR;
    }

private:
    int n_;
    bool bias_;
    float eps_;
    Tensor weight_, biasT_;
};

// ---------------------------------- GELU / QuickGELU through the reference's fp16 tables (op/CPUGELU.cpp, op/CPUQuickGELU.cpp ->
// mllm_vec_gelu_f32 / mllm_vec_gelu_quick_f32, ggml/Quantize.hpp:74-133).  The 65536-entry table is built on the host with the
// formula the reference's init_table_* uses and lives on the device; one table per kind and backend.
class CUDAGelu final : public CUDAOpBase {
public:
    CUDAGelu(Backend *bn, string name, int kind) : CUDAOpBase(bn, name), kind_(kind) {}
    ErrorCode reshape(vector<shared_ptr<Tensor>> in, vector<shared_ptr<Tensor>> out) override { same_shape(in, out); return MLLM_NO_ERROR; }
    ErrorCode execute(vector<shared_ptr<Tensor>> in, vector<shared_ptr<Tensor>> out) override {
        auto *b = cb(backend_);
        need_dense_f32(*in[0], "GELU");
        need_dense_f32(*out[0], "GELU");
        static std::map<std::pair<Backend *, int>, std::shared_ptr<Tensor>> tables;
        auto &tab = tables[{backend_, kind_}];
        if (!tab) {
            std::vector<uint16_t> t(1 << 16);
            for (int i = 0; i < (1 << 16); i++) {
                const float x = MLLM_FP16_TO_FP32((mllm_fp16_t)i);
                // gelu: the reference build (g++ -O3, FMA target) contracts 1 + (A*x)*x into one fma; tests/golden pins it
                const float g = kind_ == 0 ? 0.5f * x * (1.0f + tanhf(0.79788456080286535587989211986876f * x * std::fmaf(0.044715f * x, x, 1.0f)))
                                           : x * (1.0f / (1.0f + expf(-1.702f * x)));
                t[i] = MLLM_FP32_TO_FP16(g);
            }
            tab = std::make_shared<Tensor>(backend_);
            tab->setName(kind_ == 0 ? "gelu.table" : "quickgelu.table");
            tab->reshape(1, 1, 1, 1 << 16);
            tab->setDtype(MLLM_TYPE_F16);
            b->memory()->device_only = true;
            tab->alloc();
            b->memory()->device_only = false;
            MLLMCU_CHECK(mllmcu_memcpy_h2d(tab->rawHostPtr(), t.data(), t.size() * 2, b->stream()));
            b->sync(); // the host vector dies at scope exit
        }
        b->dropQuantTag();
        MLLMCU_CHECK(mllmcu_gelu_table(fbase(*in[0]), fbase(*out[0]), (size_t)in[0]->count(), tab->rawHostPtr(), b->stream()));
        b->kernel_launches++;
        return MLLM_NO_ERROR;
    }

private:
    int kind_;
};

// ------------------------------------------------------------------ Parameter (op/CPUParameter.cpp:17-47,83-86)
// A learned tensor loaded under the op's own name (CLIP class_embedding, ViT position embeddings); the output aliases it.
class CUDAParameter final : public CUDAOpBase {
public:
    CUDAParameter(Backend *bn, string name, int b, int h, int s, int d) : CUDAOpBase(bn, name), b_(b), h_(h), s_(s), d_(d) { weight_ = std::make_shared<Tensor>(bn); }
    ErrorCode setUp(vector<shared_ptr<Tensor>>, vector<shared_ptr<Tensor>> out) override {
        out[0]->shallowCopyFrom(weight_, false);
        return MLLM_NO_ERROR;
    }
    ErrorCode reshape(vector<shared_ptr<Tensor>>, vector<shared_ptr<Tensor>> out) override {
        if (out[0]->masterTensor() == nullptr) out[0]->shallowCopyFrom(weight_, false);
        out[0]->reshape(b_, h_, s_, d_);
        return MLLM_NO_ERROR;
    }
    ErrorCode load(AbstructLoader &loader) override {
        weight_->setName(name());
        weight_->reshape(b_, h_, s_, d_);
        const DataType dt = loader.getDataType(name());
        weight_->setDtype(dt != MLLM_TYPE_COUNT ? dt : MLLM_TYPE_F32);
        weight_->alloc();
        if (dt != MLLM_TYPE_COUNT) loader.load(weight_);
        return Op::load(loader);
    }
    ErrorCode execute(vector<shared_ptr<Tensor>>, vector<shared_ptr<Tensor>> out) override {
        if (out[0]->masterTensor() == nullptr || out[0]->masterTensor()->name() != weight_->name())
            throw std::runtime_error("[mllm-cuda] Parameter " + name() + ": the output must alias the parameter");
        return MLLM_NO_ERROR;
    }

private:
    int b_, h_, s_, d_;
    shared_ptr<Tensor> weight_;
};

// ------------------------------------------------------------------ Convolution2D (op/CPUConvolution2D.cpp:33-133, VALID)
// input [B, head = H, seq = C, dim = W] (NCHW bytes), weight name.weight [O][C][kh][kw] F32; output [B, head = oh, seq = O, dim = ow].
class CUDAConvolution2D final : public CUDAOpBase {
public:
    CUDAConvolution2D(Backend *bn, string name, const OpParam &p) :
        CUDAOpBase(bn, name), C_((int)p.at("in_channel")), O_((int)p.at("out_channel")), kh_((int)p.at("kernal_h")), kw_((int)p.at("kernal_w")),
        sh_((int)p.at("stride_h")), sw_((int)p.at("stride_w")), bias_((bool)getp(p, "bias", 1)) {
        if ((PaddingType)(int)getp(p, "padding", (float)VALID) != VALID) throw std::runtime_error("[mllm-cuda] Convolution2D " + name + ": only VALID padding (the patch-embed case)");
    }
    ErrorCode reshape(vector<shared_ptr<Tensor>> in, vector<shared_ptr<Tensor>> out) override {
        if (in[0]->sequence() != C_) throw std::runtime_error("[mllm-cuda] Convolution2D " + name() + ": channel mismatch");
        out[0]->reshape(in[0]->batch(), (in[0]->head() - kh_) / sh_ + 1, O_, (in[0]->dimension() - kw_) / sw_ + 1);
        return MLLM_NO_ERROR;
    }
    ErrorCode load(AbstructLoader &loader) override {
        load_param(weight_, backend_, name() + ".weight", O_, C_ * kh_ * kw_, loader, MLLM_TYPE_F32);
        if (bias_) load_param(biasT_, backend_, name() + ".bias", 1, O_, loader, MLLM_TYPE_F32);
        if (weight_.dtype() != MLLM_TYPE_F32) throw std::runtime_error("[mllm-cuda] Convolution2D " + name() + ": F32 kernel only");
        return Op::load(loader);
    }
    ErrorCode execute(vector<shared_ptr<Tensor>> in, vector<shared_ptr<Tensor>> out) override {
        auto *b = cb(backend_);
        Tensor &x = *in[0];
        need_dense_f32(x, "Convolution2D");
        need_dense_f32(*out[0], "Convolution2D");
        b->dropQuantTag();
        MLLMCU_CHECK(mllmcu_conv2d_valid(fbase(x), weight_.hostPtr<float>(), bias_ ? biasT_.hostPtr<float>() : nullptr, fbase(*out[0]), x.batch(), C_, x.head(),
                                         x.dimension(), O_, kh_, kw_, sh_, sw_, 0, b->stream()));
        b->kernel_launches++;
        return MLLM_NO_ERROR;
    }

private:
    int C_, O_, kh_, kw_, sh_, sw_;
    bool bias_;
    Tensor weight_, biasT_;
};

// ------------------------------------------------------------------ Convolution3D (op/CPUConvolution3D.cpp:20-120, VALID)
// input BCTHW [B][C][T][H][W], weight name.weight [O][C][kt][kh][kw] F32; output BCTHW [B][O][ot][oh][ow] (Qwen2-VL patch embed:
// one receptive field per pre-cut patch, models/qwen2_vl/modeling_qwen2_vl.hpp:27-35).
class CUDAConvolution3D final : public CUDAOpBase {
public:
    CUDAConvolution3D(Backend *bn, string name, const OpParam &p) :
        CUDAOpBase(bn, name), C_((int)p.at("in_channel")), O_((int)p.at("out_channel")), kt_((int)p.at("kernal_t")), kh_((int)p.at("kernal_h")),
        kw_((int)p.at("kernal_w")), st_((int)p.at("stride_t")), sh_((int)p.at("stride_h")), sw_((int)p.at("stride_w")), bias_((bool)getp(p, "bias", 1)) {
        if ((PaddingType)(int)getp(p, "padding", (float)VALID) != VALID) throw std::runtime_error("[mllm-cuda] Convolution3D " + name + ": only VALID padding (the patch-embed case)");
    }
    ErrorCode reshape(vector<shared_ptr<Tensor>> in, vector<shared_ptr<Tensor>> out) override {
        if (in[0]->ctype() != BCTHW || in[0]->channel() != C_) throw std::runtime_error("[mllm-cuda] Convolution3D " + name() + ": BCTHW input with in_channel channels expected");
        out[0]->reshape(in[0]->batch(), O_, (in[0]->time() - kt_) / st_ + 1, (in[0]->height() - kh_) / sh_ + 1, (in[0]->width() - kw_) / sw_ + 1);
        return MLLM_NO_ERROR;
    }
    ErrorCode load(AbstructLoader &loader) override {
        load_param(weight_, backend_, name() + ".weight", O_, C_ * kt_ * kh_ * kw_, loader, MLLM_TYPE_F32);
        if (bias_) load_param(biasT_, backend_, name() + ".bias", 1, O_, loader, MLLM_TYPE_F32);
        if (weight_.dtype() != MLLM_TYPE_F32) throw std::runtime_error("[mllm-cuda] Convolution3D " + name() + ": F32 kernel only");
        return Op::load(loader);
    }
    ErrorCode execute(vector<shared_ptr<Tensor>> in, vector<shared_ptr<Tensor>> out) override {
        auto *b = cb(backend_);
        Tensor &x = *in[0];
        if (x.dtype() != MLLM_TYPE_F32 || out[0]->dtype() != MLLM_TYPE_F32 || x.masterTensor() != nullptr) throw std::runtime_error("[mllm-cuda] Convolution3D " + name() + ": dense fp32 tensors expected");
        b->dropQuantTag();
        MLLMCU_CHECK(mllmcu_conv3d_valid(x.hostPtr<float>(), weight_.hostPtr<float>(), bias_ ? biasT_.hostPtr<float>() : nullptr, out[0]->hostPtr<float>(), x.batch(), C_,
                                         x.time(), x.height(), x.width(), O_, kt_, kh_, kw_, st_, sh_, sw_, 0, b->stream()));
        b->kernel_launches++;
        return MLLM_NO_ERROR;
    }

private:
    int C_, O_, kt_, kh_, kw_, st_, sh_, sw_;
    bool bias_;
    Tensor weight_, biasT_;
};

// ------------------------------------------------------------------ transpose (op/CPUTransposeFunc.hpp:24-205)
// The pairs swap axis LABELS one after the other, exactly as the reference swaps chls() entries (:62-70): afterwards the new
// tensor's axis L runs over what was the old tensor's axis map[L].  The reference leaves the bytes in place and re-labels them; off
// the CPU backend Tensor::transpose is never in place, so the result is written as a fresh dense BSHD tensor with one strided copy.
class CUDATranspose final : public CUDAOpBase {
public:
    CUDATranspose(Backend *bn, string name, vector<std::pair<int, int>> axes) : CUDAOpBase(bn, name) {
        for (int i = 0; i < 4; i++) map_[i] = i;
        for (auto &a : axes) {
            if (a.first < 0 || a.first > 3 || a.second < 0 || a.second > 3) throw std::runtime_error("[mllm-cuda] transpose " + name + ": 4-D axes only");
            std::swap(map_[a.first], map_[a.second]);
        }
    }
    ErrorCode reshape(vector<shared_ptr<Tensor>> in, vector<shared_ptr<Tensor>> out) override {
        const int n[4] = {in[0]->batch(), in[0]->sequence(), in[0]->head(), in[0]->dimension()}; // by Chl
        out[0]->reshape(n[map_[BATCH]], n[map_[HEAD]], n[map_[SEQUENCE]], n[map_[DIMENSION]]);
        out[0]->setDtype(in[0]->dtype());
        return MLLM_NO_ERROR;
    }
    ErrorCode execute(vector<shared_ptr<Tensor>> in, vector<shared_ptr<Tensor>> out) override {
        auto *b = cb(backend_);
        Tensor &x = *in[0], &y = *out[0];
        int xn[4], yn[4];
        long long xs[4], ys[4];
        chl_geometry(x, xn, xs);
        chl_geometry(y, yn, ys);
        // walk the output in memory order (b, s, h, d); its axis L reads the input along map_[L]
        const int order[4] = {BATCH, SEQUENCE, HEAD, DIMENSION};
        int n[4];
        long long ss[4], ds[4];
        for (int i = 0; i < 4; i++) { n[i] = yn[order[i]]; ds[i] = ys[order[i]]; ss[i] = xs[map_[order[i]]]; }
        b->dropQuantTag();
        MLLMCU_CHECK(mllmcu_copy_strided(base(x), base(y), x.dtype() == MLLM_TYPE_F16 ? 2 : 4, n, ss, ds, b->stream()));
        b->kernel_launches++;
        return MLLM_NO_ERROR;
    }

private:
    int map_[4];
};

// ------------------------------------------------------------------ flatten (op/CPUFlattenFunc.hpp:29-75): merges adjacent axes of a
// dense BSHD tensor; in memory order (b, s, h, d) every supported pair is already contiguous, so this is a reshape (run in place).
class CUDAFlatten final : public CUDAOpBase {
public:
    CUDAFlatten(Backend *bn, string name, int a0, int a1) : CUDAOpBase(bn, name), a0_(a0), a1_(a1) {}
    ErrorCode reshape(vector<shared_ptr<Tensor>> in, vector<shared_ptr<Tensor>> out) override {
        Tensor &x = *in[0];
        if (x.ctype() != BSHD) throw std::runtime_error("[mllm-cuda] flatten: BSHD tensors only");
        int b = x.batch(), h = x.head(), s = x.sequence(), d = x.dimension();
        if (a0_ == BATCH && a1_ == SEQUENCE) { if (h != 1 && b != 1) throw std::runtime_error("[mllm-cuda] flatten(BATCH, SEQUENCE): needs one head or one batch"); s *= b; b = 1; }
        else if (a0_ == HEAD && a1_ == SEQUENCE) { s *= h; h = 1; }   // new s = s * H + h: the order of the bytes
        else if (a0_ == HEAD && a1_ == DIMENSION) { d *= h; h = 1; }
        else throw std::runtime_error("[mllm-cuda] flatten: axis pair not supported");
        if (in[0].get() != out[0].get() && !dense(x)) throw std::runtime_error("[mllm-cuda] flatten: dense input expected");
        out[0]->reshape(b, h, s, d);
        return MLLM_NO_ERROR;
    }
    ErrorCode execute(vector<shared_ptr<Tensor>> in, vector<shared_ptr<Tensor>> out) override {
        if (in[0]->rawHostPtr() != out[0]->rawHostPtr()) { // not in place: the same bytes under the new shape
            auto *b = cb(backend_);
            MLLMCU_CHECK(mllmcu_memcpy_d2d(out[0]->rawHostPtr(), base(*in[0]), in[0]->cntSize(), b->stream()));
        }
        return MLLM_NO_ERROR;
    }

private:
    int a0_, a1_;
};

// ------------------------------------------------------------------ cat (op/CPUCatFunc.hpp:88-113,116-300): every input is placed with
// one strided copy; an input with batch 1 is repeated over the output's batches as the reference does (:121-127).
class CUDACat final : public CUDAOpBase {
public:
    CUDACat(Backend *bn, string name, int axis) : CUDAOpBase(bn, name), axis_(axis) {
        if (axis_ < 0 || axis_ > 3) throw std::runtime_error("[mllm-cuda] cat " + name + ": axis must be BATCH, SEQUENCE, HEAD or DIMENSION");
    }
    ErrorCode reshape(vector<shared_ptr<Tensor>> in, vector<shared_ptr<Tensor>> out) override {
        int n[4] = {0, in[0]->sequence(), in[0]->head(), in[0]->dimension()};
        for (auto &t : in) n[BATCH] = std::max(n[BATCH], t->batch());
        if (axis_ != BATCH) n[axis_] = 0;
        else n[BATCH] = 0;
        for (auto &t : in) {
            const int v[4] = {t->batch(), t->sequence(), t->head(), t->dimension()};
            for (int a = 1; a < 4; a++)
                if (a != axis_ && v[a] != n[a]) throw std::runtime_error("[mllm-cuda] cat " + name() + ": shapes differ off the concatenation axis");
            n[axis_] += v[axis_];
        }
        out[0]->reshape(n[BATCH], n[HEAD], n[SEQUENCE], n[DIMENSION]);
        out[0]->setDtype(in[0]->dtype());
        return MLLM_NO_ERROR;
    }
    ErrorCode execute(vector<shared_ptr<Tensor>> in, vector<shared_ptr<Tensor>> out) override {
        auto *b = cb(backend_);
        Tensor &y = *out[0];
        const int esz = y.dtype() == MLLM_TYPE_F16 ? 2 : 4;
        int yn[4];
        long long ys[4];
        chl_geometry(y, yn, ys);
        const int order[4] = {BATCH, SEQUENCE, HEAD, DIMENSION};
        long long at = 0;
        b->dropQuantTag();
        for (auto &t : in) {
            int xn[4];
            long long xs[4];
            chl_geometry(*t, xn, xs);
            if (t->dtype() != y.dtype()) throw std::runtime_error("[mllm-cuda] cat " + name() + ": one dtype expected");
            int n[4];
            long long ss[4], ds[4];
            for (int i = 0; i < 4; i++) { n[i] = xn[order[i]]; ss[i] = xs[order[i]]; ds[i] = ys[order[i]]; }
            if (axis_ != BATCH && xn[BATCH] == 1 && yn[BATCH] > 1) { n[0] = yn[BATCH]; ss[0] = 0; } // broadcast over batches
            MLLMCU_CHECK(mllmcu_copy_strided(base(*t), (char *)base(y) + (size_t)(at * ys[axis_]) * esz, esz, n, ss, ds, b->stream()));
            b->kernel_launches++;
            at += xn[axis_];
        }
        return MLLM_NO_ERROR;
    }

private:
    int axis_;
};

// ------------------------------------------------------------------ split (op/CPUSplitFunc.hpp:18-175) along HEAD / SEQUENCE / DIMENSION:
// the merged qkv of the Qwen2-VL vision attention (modeling_qwen2_vl.hpp:63-68).  The reference hands out child views of the input;
// here each piece is its own dense tensor (the consumers -- view, rotary, attention -- want contiguous heads).
class CUDASplit final : public CUDAOpBase {
public:
    CUDASplit(Backend *bn, string name, vector<int> each, int axis) : CUDAOpBase(bn, name), each_(std::move(each)), axis_(axis) {
        if (axis_ < 1 || axis_ > 3) throw std::runtime_error("[mllm-cuda] split " + name + ": HEAD, SEQUENCE or DIMENSION only (the D_HD / HD merged layouts are not built)");
    }
    ErrorCode reshape(vector<shared_ptr<Tensor>> in, vector<shared_ptr<Tensor>> out) override {
        if (out.size() != each_.size()) throw std::runtime_error("[mllm-cuda] split: one output per piece");
        const int n[4] = {in[0]->batch(), in[0]->sequence(), in[0]->head(), in[0]->dimension()};
        int total = 0;
        for (int e : each_) total += e;
        if (total != n[axis_]) throw std::runtime_error("[mllm-cuda] split " + name() + ": pieces do not add up to the axis length");
        for (size_t i = 0; i < out.size(); i++) {
            int m[4] = {n[0], n[1], n[2], n[3]};
            m[axis_] = each_[i];
            out[i]->reshape(m[BATCH], m[HEAD], m[SEQUENCE], m[DIMENSION]);
            out[i]->setDtype(in[0]->dtype());
        }
        return MLLM_NO_ERROR;
    }
    ErrorCode execute(vector<shared_ptr<Tensor>> in, vector<shared_ptr<Tensor>> out) override {
        auto *b = cb(backend_);
        Tensor &x = *in[0];
        const int esz = x.dtype() == MLLM_TYPE_F16 ? 2 : 4;
        int xn[4];
        long long xs[4];
        chl_geometry(x, xn, xs);
        const int order[4] = {BATCH, SEQUENCE, HEAD, DIMENSION};
        long long at = 0;
        b->dropQuantTag();
        for (auto &yt : out) {
            int yn[4];
            long long ys[4];
            chl_geometry(*yt, yn, ys);
            int n[4];
            long long ss[4], ds[4];
            for (int i = 0; i < 4; i++) { n[i] = yn[order[i]]; ss[i] = xs[order[i]]; ds[i] = ys[order[i]]; }
            MLLMCU_CHECK(mllmcu_copy_strided((const char *)base(x) + (size_t)(at * xs[axis_]) * esz, base(*yt), esz, n, ss, ds, b->stream()));
            b->kernel_launches++;
            at += yn[axis_];
        }
        return MLLM_NO_ERROR;
    }

private:
    vector<int> each_;
    int axis_;
};

// ------------------------------------------------------------------ range (op/CPURangeFunc.hpp:25-41)
class CUDARange final : public CUDAOpBase {
public:
    CUDARange(Backend *bn, string name, int start, int end) : CUDAOpBase(bn, name), start_(start), end_(end) {}
    ErrorCode reshape(vector<shared_ptr<Tensor>>, vector<shared_ptr<Tensor>> out) override {
        out[0]->reshape(1, 1, end_ - start_, 1);
        out[0]->setDtype(MLLM_TYPE_F32);
        return MLLM_NO_ERROR;
    }
    ErrorCode execute(vector<shared_ptr<Tensor>>, vector<shared_ptr<Tensor>> out) override {
        auto *b = cb(backend_);
        MLLMCU_CHECK(mllmcu_range(fbase(*out[0]), start_, end_ - start_, b->stream()));
        b->kernel_launches++;
        return MLLM_NO_ERROR;
    }

private:
    int start_, end_;
};

// ------------------------------------------------------------------ zero_like (F_LIKE, op/CPULikeFunc.hpp:26-36): the reference memsets the
// output with (int)like_value as the BYTE value, so only 0 gives what the name promises; the same bytes are written here.
class CUDALike final : public CUDAOpBase {
public:
    CUDALike(Backend *bn, string name, float v) : CUDAOpBase(bn, name), v_(v) {}
    ErrorCode reshape(vector<shared_ptr<Tensor>> in, vector<shared_ptr<Tensor>> out) override {
        same_shape(in, out);
        out[0]->setDtype(in[0]->dtype());
        return MLLM_NO_ERROR;
    }
    ErrorCode execute(vector<shared_ptr<Tensor>>, vector<shared_ptr<Tensor>> out) override {
        auto *b = cb(backend_);
        MLLMCU_CHECK(mllmcu_memset(out[0]->rawHostPtr(), (int)v_, out[0]->cntSize(), b->stream()));
        return MLLM_NO_ERROR;
    }

private:
    float v_;
};

// ------------------------------------------------------------------ where (op/CPUWhereFunc.hpp:27-95): positions with x == value in the
// reference's visit order, compacted on the device; the host reads ONE integer (the count) because the output's shape depends on
// the data -- the reference allocates its output inside execute for the same reason.
class CUDAWhere final : public CUDAOpBase {
public:
    CUDAWhere(Backend *bn, string name, float value, int axis) : CUDAOpBase(bn, name), value_(value), axis_(axis) {}
    ErrorCode reshape(vector<shared_ptr<Tensor>>, vector<shared_ptr<Tensor>>) override { return MLLM_NO_ERROR; }
    ErrorCode setUp(vector<shared_ptr<Tensor>>, vector<shared_ptr<Tensor>>) override { return MLLM_NO_ERROR; }
    ErrorCode execute(vector<shared_ptr<Tensor>> in, vector<shared_ptr<Tensor>> out) override {
        auto *b = cb(backend_);
        Tensor &x = *in[0];
        need_dense_f32(x, "where");
        const long long total = x.count();
        // scratch: [count][linear indices]; managed memory, so the count is readable after the sync
        if (!scratch_ || scratch_->count() < total + 1) {
            scratch_ = std::make_shared<Tensor>(backend_);
            scratch_->setName(name() + ".scratch");
            scratch_->reshape(1, 1, 1, (int)total + 1);
            scratch_->setDtype(MLLM_TYPE_F32); // 4-byte slots holding int32
            scratch_->alloc();
        }
        int *cnt = (int *)scratch_->rawHostPtr(), *lin = cnt + 1;
        MLLMCU_CHECK(mllmcu_where_eq(fbase(x), total, value_, lin, cnt, b->stream()));
        b->sync();
        const int num = *cnt;
        const int n[4] = {x.batch(), x.sequence(), x.head(), x.dimension()}; // visit order b, s, h, d
        Tensor &y = *out[0];
        y.setDtype(MLLM_TYPE_F32);
        if (axis_ == -1) {
            y.reshape(1, 1, 4, num);
            y.alloc();
            const int axis_of_row[4] = {0, 2, 1, 3}; // rows b, h, s, d (CPUWhereFunc.hpp:61-76)
            for (int r = 0; r < 4 && num > 0; r++) MLLMCU_CHECK(mllmcu_where_unravel(lin, num, n, axis_of_row[r], fbase(y) + (size_t)r * num, b->stream()));
        } else {
            y.reshape(1, 1, 1, num);
            y.alloc();
            const int a = axis_ == BATCH ? 0 : axis_ == SEQUENCE ? 1 : axis_ == HEAD ? 2 : 3;
            if (num > 0) MLLMCU_CHECK(mllmcu_where_unravel(lin, num, n, a, fbase(y), b->stream()));
        }
        b->kernel_launches += 2;
        return MLLM_NO_ERROR;
    }

private:
    float value_;
    int axis_;
    shared_ptr<Tensor> scratch_;
};

// ------------------------------------------------------------------ index_put (op/CPUIndexPutFunc.hpp:27-131)
//   accumulate == false (Qwen2-VL): rows of the destination at the given sequence positions are overwritten in place;
//   accumulate == true  (LLaVA): the ONE position given is replaced by all rows of the source (output grows by R - 1 rows).
// The indices are read by the kernels; the host never sees them.
class CUDAIndexPut final : public CUDAOpBase {
public:
    CUDAIndexPut(Backend *bn, string name, bool accumulate) : CUDAOpBase(bn, name), acc_(accumulate) {}
    ErrorCode setUp(vector<shared_ptr<Tensor>>, vector<shared_ptr<Tensor>>) override { return MLLM_NO_ERROR; }
    ErrorCode reshape(vector<shared_ptr<Tensor>> in, vector<shared_ptr<Tensor>> out) override {
        Tensor &dest = *in[0];
        if (in.size() > 1 && in[1]->batch() == 0) { // no image: the text embeddings pass through
            if (acc_) throw std::runtime_error("[mllm-cuda] index_put " + name() + ": empty source with accumulate");
            out[0]->reshape(dest.batch(), 1, dest.sequence(), dest.dimension());
            return MLLM_NO_ERROR;
        }
        if (in.size() != 3 || dest.batch() != 1 || dest.head() != 1 || in[1]->head() != 1 || dest.dimension() != in[1]->dimension())
            throw std::runtime_error("[mllm-cuda] index_put " + name() + ": {dest [1,1,S,D], src [n,1,R,D], idx} expected");
        if (!acc_) out[0]->reshape(1, 1, dest.sequence(), dest.dimension());
        else {
            if (in[1]->batch() != 1) throw std::runtime_error("[mllm-cuda] index_put " + name() + ": one spliced source (one image) per call");
            out[0]->reshape(1, 1, dest.sequence() - 1 + in[1]->sequence(), dest.dimension());
        }
        out[0]->setDtype(dest.dtype());
        return MLLM_NO_ERROR;
    }
    ErrorCode execute(vector<shared_ptr<Tensor>> in, vector<shared_ptr<Tensor>> out) override {
        if (in.size() > 1 && in[1]->batch() == 0) return MLLM_NO_ERROR;
        auto *b = cb(backend_);
        Tensor &dest = *in[0], &src = *in[1], &idx = *in[2];
        need_dense_f32(dest, "index_put");
        need_dense_f32(src, "index_put");
        need_dense_f32(idx, "index_put");
        if (idx.batch() != 1 || idx.sequence() != 1 || idx.head() != 1) throw std::runtime_error("[mllm-cuda] index_put " + name() + ": indices must be [1,1,1,n]");
        b->dropQuantTag();
        if (!acc_) {
            if (idx.dimension() > src.sequence()) throw std::runtime_error("[mllm-cuda] index_put " + name() + ": more positions than source rows");
            MLLMCU_CHECK(mllmcu_index_put_rows(fbase(dest), dest.dimension(), fbase(src), src.dimension(), fbase(idx), idx.dimension(), dest.dimension(), b->stream()));
            if (base(*out[0]) != base(dest)) // not in place: the result is the updated destination
                MLLMCU_CHECK(mllmcu_memcpy_d2d(base(*out[0]), base(dest), dest.cntSize(), b->stream()));
        } else {
            if (idx.dimension() != 1) throw std::runtime_error("[mllm-cuda] index_put " + name() + ": exactly one image position expected (found " + std::to_string(idx.dimension()) + ")");
            need_dense_f32(*out[0], "index_put");
            MLLMCU_CHECK(mllmcu_index_splice(fbase(dest), fbase(src), fbase(idx), fbase(*out[0]), dest.sequence(), src.sequence(), dest.dimension(), b->stream()));
        }
        b->kernel_launches++;
        return MLLM_NO_ERROR;
    }

private:
    bool acc_;
};

// ------------------------------------------------------------------ VisionRoPE (op/CPUVisionRoPE.cpp:18-215): the [tokens][2 * (dim/2)]
// table of rotary angles of the Qwen2-VL vision tower, from the (t, h, w) grid: token order = merge-window order (rot_pos_emb,
// :52-91), each row = [h position * inv_freq | w position * inv_freq].  A position table, built on the host like the reference's
// (and like the RoPE sin / cos tables of this backend) and uploaded; the grid is three numbers of an input tensor.
class CUDAVisionRoPE final : public CUDAOpBase {
public:
    CUDAVisionRoPE(Backend *bn, string name, int dim, int merge) : CUDAOpBase(bn, name), dim_(dim), merge_(merge) {}
    ErrorCode reshape(vector<shared_ptr<Tensor>> in, vector<shared_ptr<Tensor>> out) override {
        read_grid(*in[0]);
        out[0]->reshape(1, 1, t_ * h_ * w_, 2 * (dim_ / 2));
        return MLLM_NO_ERROR;
    }
    ErrorCode execute(vector<shared_ptr<Tensor>> in, vector<shared_ptr<Tensor>> out) override {
        auto *b = cb(backend_);
        read_grid(*in[0]);
        const int half = dim_ / 2, cols = 2 * half;
        std::vector<float> inv(half);
        for (int i = 0; i < half; i++) inv[i] = 1.0f / std::pow(10000.0f, (2.0f * i) / (float)dim_);
        std::vector<float> tab((size_t)t_ * h_ * w_ * cols);
        const int nh = h_ / merge_, nw = w_ / merge_;
        size_t row = 0;
        for (int t = 0; t < t_; t++)
            for (int blk = 0; blk < nh * nw; blk++)
                for (int jh = 0; jh < merge_; jh++)
                    for (int jw = 0; jw < merge_; jw++, row++) {
                        const float hp = (float)((blk / nw) * merge_ + jh), wp = (float)((blk % nw) * merge_ + jw);
                        for (int i = 0; i < half; i++) { tab[row * cols + i] = hp * inv[i]; tab[row * cols + half + i] = wp * inv[i]; }
                    }
        need_dense_f32(*out[0], "VisionRoPE");
        MLLMCU_CHECK(mllmcu_memcpy_h2d(fbase(*out[0]), tab.data(), tab.size() * 4, b->stream()));
        b->sync(); // the host vector dies at scope exit
        return MLLM_NO_ERROR;
    }

private:
    void read_grid(Tensor &g) { // three floats of a managed tensor, read on the host as the reference reads them
        cb(backend_)->sync();
        t_ = (int)g.dataAt<float>(0, 0, 0, 0); h_ = (int)g.dataAt<float>(0, 0, 0, 1); w_ = (int)g.dataAt<float>(0, 0, 0, 2);
        if (t_ <= 0 || h_ <= 0 || w_ <= 0 || h_ % merge_ || w_ % merge_) throw std::runtime_error("[mllm-cuda] VisionRoPE " + name() + ": bad grid");
    }
    int dim_, merge_, t_ = 0, h_ = 0, w_ = 0;
};

// Tensor::apply_rotary_pos_emb_vision (F_APPLY_VISIOROPE, op/CPUVisionRoPEFunc.hpp:20-160): x [B,H,S,D], rotary_pos_emb [1,1,S,D/2];
// sin / cos of the angles are evaluated on the device (<= 2 ulp from libm, tests/test_refops_gpu.py states the tolerance).
class CUDAApplyVisionRoPE final : public CUDAOpBase {
public:
    using CUDAOpBase::CUDAOpBase;
    ErrorCode reshape(vector<shared_ptr<Tensor>> in, vector<shared_ptr<Tensor>> out) override { same_shape(in, out); return MLLM_NO_ERROR; }
    ErrorCode execute(vector<shared_ptr<Tensor>> in, vector<shared_ptr<Tensor>> out) override {
        auto *b = cb(backend_);
        Tensor &x = *in[0];
        need_dense_f32(x, "apply_rotary_pos_emb_vision");
        need_dense_f32(*in[1], "apply_rotary_pos_emb_vision");
        need_dense_f32(*out[0], "apply_rotary_pos_emb_vision");
        if (in[1]->dimension() * 2 != x.dimension() || in[1]->sequence() != x.sequence()) throw std::runtime_error("[mllm-cuda] vision rope: rotary_pos_emb must be [1,1,S,D/2]");
        b->dropQuantTag();
        MLLMCU_CHECK(mllmcu_rope_vision(fbase(x), fbase(*in[1]), nullptr, fbase(*out[0]), x.batch(), x.sequence(), x.head(), x.dimension(), 0, b->stream()));
        b->kernel_launches++;
        return MLLM_NO_ERROR;
    }
};

// ------------------------------------------------------------------ MultimodalRoPE (op/CPUMultimodalRoPE.cpp:25-331)
// inputs {x [B,H,S,D], position_ids [3,1,1,S]}.  The sin / cos tables ([S][D], the t / h / w sections gathered) are rebuilt on the
// HOST from the ids exactly as multimodal_sinusoidal_position_embedding + apply_multimodal_rotary_pos_emb do (same libm sinf /
// cosf), so the rotation is bit-identical to the reference; q and k of one layer -- and all layers of one step -- see the same
// ids, so the tables are cached on the ids' bytes and rebuilt once per step.
class CUDAMultimodalRoPE final : public CUDAOpBase {
public:
    CUDAMultimodalRoPE(Backend *bn, string name, const OpParam &p) : CUDAOpBase(bn, name), theta_(p.at("rope_theta")) {
        for (int i = 0;; i++) {
            auto it = p.find("mrope_section_" + std::to_string(i));
            if (it == p.end()) break;
            sec_.push_back((int)it->second);
        }
    }
    ErrorCode reshape(vector<shared_ptr<Tensor>> in, vector<shared_ptr<Tensor>> out) override { same_shape(in, out); return MLLM_NO_ERROR; }
    ErrorCode execute(vector<shared_ptr<Tensor>> in, vector<shared_ptr<Tensor>> out) override {
        auto *b = cb(backend_);
        Tensor &x = *in[0], &pid = *in[1], &y = *out[0];
        need_dense_f32(x, "MultimodalRoPE");
        const int S = x.sequence(), D = x.dimension(), half = D / 2;
        if (pid.batch() != 3 || pid.dimension() < S || pid.dtype() != MLLM_TYPE_F32) throw std::runtime_error("[mllm-cuda] MultimodalRoPE " + name() + ": position_ids must be [3,1,1,S] fp32");
        struct Tables { std::vector<float> ids; float theta = 0; int S = 0, D = 0; std::vector<int> sec; shared_ptr<Tensor> sin_t, cos_t; };
        static std::map<Backend *, Tables> cache;
        Tables &tb = cache[backend_];
        b->sync(); // the ids are a managed tensor written by the host or by an upload
        std::vector<float> pos((size_t)3 * S);
        for (int a = 0; a < 3; a++)
            for (int s = 0; s < S; s++) pos[(size_t)a * S + s] = pid.dataAt<float>(a, 0, 0, s);
        if (!(tb.sin_t && tb.S == S && tb.D == D && tb.theta == theta_ && tb.sec == sec_ && tb.ids == pos)) {
            std::vector<float> theta(half), sn((size_t)S * D, 0.f), cs((size_t)S * D, 0.f);
            for (int i = 0; i < half; i++) theta[i] = (float)(1.0 / pow((double)theta_, 2.0 * i / D));
            int start = 0;
            for (size_t j = 0; j < sec_.size(); j++) {
                const int axis = (int)(j % 3);
                for (int c = start; c < start + sec_[j] && c < D; c++)
                    for (int s = 0; s < S; s++) {
                        const float value = theta[c % half] * pos[(size_t)axis * S + s];
                        cs[(size_t)s * D + c] = cosf(value) * 1.0f;
                        sn[(size_t)s * D + c] = sinf(value) * 1.0f;
                    }
                start += sec_[j];
            }
            auto mk = [&](const char *nm) {
                auto t = std::make_shared<Tensor>(backend_);
                t->setName(nm);
                t->reshape(1, 1, S, D);
                t->setDtype(MLLM_TYPE_F32);
                t->alloc();
                return t;
            };
            tb.sin_t = mk("mrope.sin");
            tb.cos_t = mk("mrope.cos");
            MLLMCU_CHECK(mllmcu_memcpy_h2d(tb.sin_t->rawHostPtr(), sn.data(), sn.size() * 4, b->stream()));
            MLLMCU_CHECK(mllmcu_memcpy_h2d(tb.cos_t->rawHostPtr(), cs.data(), cs.size() * 4, b->stream()));
            b->sync(); // the host vectors die at scope exit
            tb.ids = pos; tb.theta = theta_; tb.S = S; tb.D = D; tb.sec = sec_;
        }
        const View ys = strides(y);
        const int f16 = y.dtype() == MLLM_TYPE_F16;
        b->dropQuantTag();
        MLLMCU_CHECK(mllmcu_rope_mrope(fbase(x), base(y), x.batch(), S, x.head(), D, D, tb.sin_t->hostPtr<float>(), tb.cos_t->hostPtr<float>(), D, f16, ys.bs, ys.ss, b->stream()));
        b->kernel_launches++;
        return MLLM_NO_ERROR;
    }

    // decode plan: a text token has the same position on the t / h / w axes, so its table row is the rotate-half (HF) row of that
    // position when the sections cover the D / 2 frequencies; the plan builds that table from theta (CUDADecodePlan.cpp)
    bool fuseInfo(CUDAFuseInfo &f) override {
        int sum = 0;
        for (int v : sec_) sum += v;
        f.mrope = 1; f.rope_type = HFHUBROPE; f.rot_dim = 2 * sum; f.rope_theta = theta_;
        return sum > 0;
    }

private:
    float theta_;
    vector<int> sec_;
};

// ------------------------------------------------------------------ creators
#define CREATOR(NAME, EXPR)                                                                                   \
    class NAME final : public CUDABackend::Creator {                                                          \
    public:                                                                                                   \
        Op *create(OpParam p, Backend *bn, string name, int threadCount) const override { return EXPR; }     \
    }
CREATOR(LayerNormCreator, new CUDALayerNorm(bn, name, (int)p.at("norm_size"), (bool)getp(p, "bias", 1), getp(p, "epsilon", 1e-6f)));
CREATOR(GeluCreator, new CUDAGelu(bn, name, 0));
CREATOR(QuickGeluCreator, new CUDAGelu(bn, name, 1));
CREATOR(ParameterCreator, new CUDAParameter(bn, name, (int)p.at("batch"), (int)p.at("head"), (int)p.at("seq"), (int)p.at("dim")));
CREATOR(Conv2DCreator, new CUDAConvolution2D(bn, name, p));
CREATOR(Conv3DCreator, new CUDAConvolution3D(bn, name, p));
static Op *make_transpose(const OpParam &p, Backend *bn, const string &name) { // Tensor::transpose packs the pairs, Tensor.cpp:514-522
    vector<std::pair<int, int>> axes;
    for (int i = 0; i < (int)p.at("num_pairs"); i++) axes.push_back({(int)p.at("axis1_" + std::to_string(i)), (int)p.at("axis2_" + std::to_string(i))});
    return new CUDATranspose(bn, name, axes);
}
CREATOR(TransposeCreator, make_transpose(p, bn, name));
CREATOR(FlattenCreator, new CUDAFlatten(bn, name, (int)p.at("axis_start"), (int)p.at("axis_end")));
CREATOR(CatCreator, new CUDACat(bn, name, (int)p.at("axis")));
static Op *make_split(const OpParam &p, Backend *bn, const string &name) {
    vector<int> each;
    for (int i = 0; i < (int)p.at("num_splits"); i++) each.push_back((int)p.at("dim_" + std::to_string(i)));
    return new CUDASplit(bn, name, each, (int)p.at("split_dim"));
}
CREATOR(SplitCreator, make_split(p, bn, name));
CREATOR(RangeCreator, new CUDARange(bn, name, (int)p.at("start"), (int)p.at("end")));
CREATOR(LikeCreator, new CUDALike(bn, name, p.at("like_value")));
CREATOR(WhereCreator, new CUDAWhere(bn, name, p.at("value"), (int)p.at("axis")));
CREATOR(IndexPutCreator, new CUDAIndexPut(bn, name, (bool)p.at("accumulate")));
CREATOR(VisionRoPECreator, new CUDAVisionRoPE(bn, name, (int)p.at("dim"), (int)p.at("spatial_merge_size")));
CREATOR(ApplyVisionRoPECreator, new CUDAApplyVisionRoPE(bn, name));
CREATOR(MultimodalRoPECreator, new CUDAMultimodalRoPE(bn, name, p));

void CUDABackend::registerVisionOps() {
    addCreator(LAYERNORM, new LayerNormCreator());
    addCreator(OP_GELU, new GeluCreator());
    addCreator(QUICKGLUE, new QuickGeluCreator());
    addCreator(PARAMETER, new ParameterCreator());
    addCreator(CONVOLUTION2D, new Conv2DCreator());
    addCreator(CONVOLUTION3D, new Conv3DCreator());
    addCreator(F_TRANPOSE, new TransposeCreator());
    addCreator(F_FLATTEN, new FlattenCreator());
    addCreator(F_CAT, new CatCreator());
    addCreator(F_SPLIT, new SplitCreator());
    addCreator(F_RANGE, new RangeCreator());
    addCreator(F_WHERE, new WhereCreator());
    addCreator(F_LIKE, new LikeCreator());
    addCreator(F_INDEX_PUT, new IndexPutCreator());
    addCreator(VISIONROPE, new VisionRoPECreator());
    addCreator(F_APPLY_VISIOROPE, new ApplyVisionRoPECreator());
    addCreator(MULTIMODALROPE, new MultimodalRoPECreator());
}

} // namespace mllm
