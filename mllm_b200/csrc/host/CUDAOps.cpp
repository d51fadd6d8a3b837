// CUDAOps.cpp -- the CUDA twins of the CPU op classes on the hot path.  Each class mirrors
// its reference op's constructor parameters (OpParam keys set by mllm/Layer.hpp), weight
// names, reshape rules and execute semantics, and forwards the arithmetic to one C-ABI
// call (include/mllmcu.h).  Reference op -> class:
//   CPULinear (op/CPULinear.cpp:22-233)            -> CUDALinear
//   CPURMSNorm (op/CPURMSNorm.cpp)                 -> CUDARMSNorm
//   CPULayerNorm (op/CPULayerNorm.cpp)             -> CUDALayerNorm
//   CPURoPE (op/CPURoPE.cpp)                       -> CUDARoPE
//   CPUKVCache (op/CPUKVCache.cpp)                 -> CUDAKVCache
//   CPUEmbedding (op/CPUEmbedding.cpp)             -> CUDAEmbedding
//   CPUSiLU (op/CPUSiLU.cpp)                       -> CUDASiLU
//   CPUaddTwoFunction / CPUmulTwoFunction (op/CPUBinaryFunc.hpp) -> CUDABinary
//   CPUviewFunction (op/CPUViewFunc.hpp)           -> CUDAView
//   CPUclipFunction (op/CPUClipFunc.hpp)           -> CUDAClip
//   CPUFlashAttention2Func (op/CPUFlashAttention2Func.hpp) -> CUDAFlashAttention2
//   CPUGELU / CPUQuickGELU (op/CPUGELU.cpp, op/CPUQuickGELU.cpp) -> CUDAGelu
//   CPUSoftMax (op/CPUSoftMax.cpp)                 -> CUDASoftMax
//   CPUmmFunction (op/CPUMatmulFunc.hpp)           -> CUDAMatmulFunc
//   CPUdivintFunction (op/CPUBinaryFunc.hpp)       -> CUDADivScalar
//   CPUtransposeFunction (op/CPUTransposeFunc.hpp) -> CUDATranspose (HEAD <-> SEQUENCE, physical)
//   CPUConvolution2D (op/CPUConvolution2D.cpp)     -> CUDAConvolution2D
//   CPUMultimodalRoPE (op/CPUMultimodalRoPE.cpp)   -> CUDAMultimodalRoPE
//   CPUVisionRoPEFuncFunction (op/CPUVisionRoPEFunc.hpp) -> CUDAVisionRoPEFunc
#include "mllm_cuda.hpp"

#include <cmath>
#include <cstring>
#include <iostream>

namespace mllm {

static CUDABackend *cb(Backend *b) { return static_cast<CUDABackend *>(b); }
static float getp(const OpParam &p, const char *k, float dflt) {
    auto it = p.find(k);
    return it == p.end() ? dflt : it->second;
}

// ------------------------------------------------------------------ Linear
class CUDALinear final : public Op {
public:
    CUDALinear(Backend *bn, string name, int in_features, int out_features, bool bias)
        : Op(bn, name), in_(in_features), out_(out_features), bias_(bias) {}
    ErrorCode reshape(vector<shared_ptr<Tensor>> in, vector<shared_ptr<Tensor>> out) override {
        if (in[0]->head() != 1 || in[0]->dimension() != in_) throw std::runtime_error("[mllm-cuda] Linear " + name() + ": input must be [B,1,S," + std::to_string(in_) + "]");
        out[0]->reshape(in[0]->batch(), in[0]->head(), in[0]->sequence(), out_);
        return MLLM_NO_ERROR;
    }
    ErrorCode load(ParamLoader &loader) override { // CPULinear.cpp:47-95
        weight_ = Tensor(backend_);
        weight_.setName(name() + ".weight");
        weight_.reshape(1, 1, out_, in_);
        DataType dt = loader.getDataType(weight_.name());
        const bool absent = (dt == MLLM_TYPE_COUNT);
        if (absent) dt = Op::noLoadWeightsDtype();
        if (dt != MLLM_TYPE_F32 && dt != MLLM_TYPE_Q4_0 && dt != MLLM_TYPE_Q8_0 && dt != MLLM_TYPE_Q4_K && dt != MLLM_TYPE_Q6_K && dt != MLLM_TYPE_Q4_0_4_4)
            throw std::runtime_error("[mllm-cuda] Linear " + name() + ": weight dtype " + std::to_string((int)dt) + " not supported by the CUDA backend yet");
        if ((dt == MLLM_TYPE_Q4_K || dt == MLLM_TYPE_Q6_K) && in_ % 256) throw std::runtime_error("[mllm-cuda] Linear " + name() + ": k-quant weights need in_features % 256 == 0");
        if (dt == MLLM_TYPE_Q4_0_4_4) {
            // shipped *-q4_0_4_4.mllm: upload the 4-row-interleaved blocks, de-interleave once on the device into plain Q4_0 rows
            // (same nibbles and scales; ggml/GemmPack.cpp:45-60), after which this is an ordinary Q4_0 Linear
            if (out_ % 4) throw std::runtime_error("[mllm-cuda] Linear " + name() + ": Q4_0_4_4 needs out_features % 4 == 0");
            Tensor packed(backend_);
            packed.setName(weight_.name());
            packed.reshape(1, 1, out_, in_);
            packed.setDtype(MLLM_TYPE_Q4_0_4_4);
            packed.alloc();
            loader.load(&packed);
            weight_.setDtype(MLLM_TYPE_Q4_0);
            weight_.alloc();
            MLLMCU_CHECK(mllmcu_q4_0_4x4_to_q4_0(packed.rawDevPtr(), weight_.rawDevPtr(), out_, in_, cb(backend_)->stream()));
            cb(backend_)->sync();
        } else {
            weight_.setDtype(dt);
            weight_.alloc();
            if (!absent) loader.load(&weight_);
        }
        if (bias_) {
            biasT_ = Tensor(backend_);
            biasT_.setName(name() + ".bias");
            biasT_.reshape(1, 1, 1, out_);
            biasT_.setDtype(MLLM_TYPE_F32);
            biasT_.alloc();
            if (loader.getDataType(biasT_.name()) != MLLM_TYPE_COUNT) loader.load(&biasT_);
        }
        return MLLM_NO_ERROR;
    }
    ErrorCode execute(vector<shared_ptr<Tensor>> in, vector<shared_ptr<Tensor>> out) override {
        auto *b = cb(backend_);
        Tensor &x = *in[0], &y = *out[0];
        const int f16 = y.dtype() == MLLM_TYPE_F16;
        const float *bias = bias_ ? biasT_.devPtr<float>() : nullptr;
        const size_t esz = f16 ? 2 : 4;
        if (weight_.dtype() == MLLM_TYPE_F32) { // weights the quantizer keeps F32 (patch embeds, small heads): fp32 GEMM, vec_dot_fp32's products in another order
            if (f16 || !x.contiguous() || !y.contiguous()) throw std::runtime_error("[mllm-cuda] Linear " + name() + ": F32 weights need dense fp32 activations");
            const int M = x.batch() * x.sequence();
            MLLMCU_CHECK(mllmcu_sgemm(x.devPtr<float>(), in_, 0, weight_.devPtr<float>(), in_, 1, 0, bias, y.devPtr<float>(), out_, 0, M, out_, in_, 1, b->stream()));
            b->dropQuantTag();
            b->kernel_launches++;
            return MLLM_NO_ERROR;
        }
        // one call when both sides are dense over (batch, seq); else one call per batch row-block
        const bool dense = x.contiguous() && (y.batch() == 1 || y.batchStride() == (long long)y.sequence() * y.seqStride());
        const int nb = dense ? 1 : x.batch();
        const int M = dense ? x.batch() * x.sequence() : x.sequence();
        const size_t wsb = mllmcu_linear_workspace_bytes(M, in_);
        void *ws = b->linearWorkspace(wsb);
        const int kind = nb == 1 ? mllmcu_linear_quant_kind((int)weight_.dtype(), M, out_, in_, weight_.rawDevPtr()) : 0;
        for (int i = 0; i < nb; i++) {
            const float *xp = x.devPtr<float>() + (size_t)i * x.batchStride();
            char *yp = (char *)y.rawDevPtr() + (size_t)i * y.batchStride() * esz;
            // q/k/v and gate/up read the same tensor: the second and third Linear reuse the first one's quantised image
            const int reuse = b->quantTagMatches(xp, M, in_, (int)x.seqStride(), kind) ? 1 : 0;
            MLLMCU_CHECK(mllmcu_linear_ex((int)weight_.dtype(), weight_.rawDevPtr(), xp, bias, yp, M, out_, in_, (int)x.seqStride(),
                                          (int)y.seqStride(), f16, ws, b->workspaceBytes(), reuse, b->stream()));
            if (kind) b->setQuantTag(xp, M, in_, (int)x.seqStride(), kind);
            else b->dropQuantTag();
            b->kernel_launches += 2 - reuse;
        }
        return MLLM_NO_ERROR;
    }
    ErrorCode free(vector<shared_ptr<Tensor>>, vector<shared_ptr<Tensor>>) override { weight_.free(); biasT_.free(); return MLLM_NO_ERROR; }
    bool fuseInfo(FuseInfo &f) override {
        if (!weight_.allocated()) return false;
        f.weight = weight_.rawDevPtr(); f.wtype = (int)weight_.dtype(); f.in = in_; f.out = out_;
        f.bias = bias_ ? biasT_.devPtr<float>() : nullptr;
        return true;
    }

private:
    int in_, out_;
    bool bias_;
    Tensor weight_, biasT_;
};

// ----------------------------------------------------------------- RMSNorm
class CUDARMSNorm final : public Op {
public:
    CUDARMSNorm(Backend *bn, string name, int norm_size, float eps, bool add_unit_offset)
        : Op(bn, name), n_(norm_size), eps_(eps), unit_(add_unit_offset) {}
    ErrorCode reshape(vector<shared_ptr<Tensor>> in, vector<shared_ptr<Tensor>> out) override {
        out[0]->reshape(in[0]->batch(), in[0]->head(), in[0]->sequence(), in[0]->dimension());
        return MLLM_NO_ERROR;
    }
    ErrorCode load(ParamLoader &loader) override {
        weight_ = Tensor(backend_);
        weight_.setName(name() + ".weight");
        weight_.reshape(1, 1, 1, n_);
        DataType dt = loader.getDataType(weight_.name());
        if (dt != MLLM_TYPE_COUNT && dt != MLLM_TYPE_F32) throw std::runtime_error("[mllm-cuda] RMSNorm " + name() + ": only F32 weights (quantizer keeps *norm* F32)");
        weight_.setDtype(MLLM_TYPE_F32);
        weight_.alloc();
        if (dt != MLLM_TYPE_COUNT) loader.load(&weight_);
        return MLLM_NO_ERROR;
    }
    ErrorCode execute(vector<shared_ptr<Tensor>> in, vector<shared_ptr<Tensor>> out) override {
        auto *b = cb(backend_);
        const int rows = in[0]->batch() * in[0]->sequence() * in[0]->head();
        MLLMCU_CHECK(mllmcu_rmsnorm(in[0]->devPtr<float>(), weight_.devPtr<float>(), out[0]->devPtr<float>(), rows, in[0]->dimension(), eps_, unit_, b->stream()));
        b->kernel_launches++;
        return MLLM_NO_ERROR;
    }

    bool fuseInfo(FuseInfo &f) override {
        if (!weight_.allocated()) return false;
        f.norm_w = weight_.devPtr<float>(); f.eps = eps_; f.unit_offset = unit_; f.in = n_;
        return true;
    }

private:
    int n_;
    float eps_;
    bool unit_;
    Tensor weight_;
};

class CUDALayerNorm final : public Op {
public:
    CUDALayerNorm(Backend *bn, string name, int norm_size, bool bias, float eps) : Op(bn, name), n_(norm_size), bias_(bias), eps_(eps) {}
    ErrorCode reshape(vector<shared_ptr<Tensor>> in, vector<shared_ptr<Tensor>> out) override {
        out[0]->reshape(in[0]->batch(), in[0]->head(), in[0]->sequence(), in[0]->dimension());
        return MLLM_NO_ERROR;
    }
    ErrorCode load(ParamLoader &loader) override {
        auto mk = [&](Tensor &t, const string &nm) {
            t = Tensor(backend_);
            t.setName(nm);
            t.reshape(1, 1, 1, n_);
            t.setDtype(MLLM_TYPE_F32);
            t.alloc();
            if (loader.getDataType(nm) != MLLM_TYPE_COUNT) loader.load(&t);
        };
        mk(weight_, name() + ".weight");
        if (bias_) mk(biasT_, name() + ".bias");
        return MLLM_NO_ERROR;
    }
    ErrorCode execute(vector<shared_ptr<Tensor>> in, vector<shared_ptr<Tensor>> out) override {
        auto *b = cb(backend_);
        const int rows = in[0]->batch() * in[0]->sequence() * in[0]->head();
        MLLMCU_CHECK(mllmcu_layernorm(in[0]->devPtr<float>(), weight_.devPtr<float>(), bias_ ? biasT_.devPtr<float>() : nullptr,
                                      out[0]->devPtr<float>(), rows, in[0]->dimension(), eps_, b->stream()));
        b->kernel_launches++;
        return MLLM_NO_ERROR;
    }

private:
    int n_;
    bool bias_;
    float eps_;
    Tensor weight_, biasT_;
};

// -------------------------------------------------------------------- RoPE
// Tables are built on the host exactly as the reference builds them (CPURoPE.cpp:23-128:
// theta in double->float, t = s*theta in float, libm sin/cos) and uploaded once per
// (pose_type, dim, theta, pos_max); like the reference's static sin_/cos_ they are shared.
struct RopeTables { Tensor sin_t, cos_t; int dim = 0, pos_max = 0; };
static map<string, shared_ptr<RopeTables>> g_rope_tables;

class CUDARoPE final : public Op {
public:
    CUDARoPE(Backend *bn, string name, const OpParam &p) : Op(bn, name) {
        pose_type_ = (int)getp(p, "pose_type", LLAMAROPE);
        rope_theta_ = getp(p, "rope_theta", 10000.f);
        partial_ = getp(p, "partial_rotary_factor", 1.0f);
        pos_max_ = (int)getp(p, "max_position_embeddings", 16384);
        if (pose_type_ != LLAMAROPE && pose_type_ != HFHUBROPE) throw std::runtime_error("[mllm-cuda] RoPE " + name + ": pose_type not supported");
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
        char key[256];
        snprintf(key, sizeof key, "%p-%d-%d-%.9g-%d-%d-%g-%g-%g-%g", (void *)backend_, pose_type_, ishape_, rope_theta_, pos_max_, theta_type_, l3_factor_, l3_low_, l3_high_, l3_old_ctx_);
        auto it = g_rope_tables.find(key);
        if (it == g_rope_tables.end()) {
            auto t = std::make_shared<RopeTables>();
            build(*t);
            g_rope_tables[key] = t;
            tab_ = t;
        } else {
            tab_ = it->second;
        }
        return MLLM_NO_ERROR;
    }
    ErrorCode execute(vector<shared_ptr<Tensor>> in, vector<shared_ptr<Tensor>> out) override {
        auto *b = cb(backend_);
        Tensor &x = *in[0], &y = *out[0];
        if (h_cnt_ + x.sequence() > pos_max_) throw std::runtime_error("[mllm-cuda] RoPE " + name() + ": position exceeds max_position_embeddings");
        MLLMCU_CHECK(mllmcu_rope(pose_type_, x.devPtr<float>(), y.rawDevPtr(), x.batch(), x.sequence(), x.head(), x.dimension(), ishape_, h_cnt_,
                                 tab_->sin_t.devPtr<float>(), tab_->cos_t.devPtr<float>(), ishape_, y.dtype() == MLLM_TYPE_F16,
                                 y.batchStride(), y.seqStride(), b->stream()));
        b->kernel_launches++;
        h_cnt_ += x.sequence(); // CPURoPE.cpp:449-515
        return MLLM_NO_ERROR;
    }
    void clearCache() override { h_cnt_ = 0; }
    bool fuseInfo(FuseInfo &f) override {
        if (!tab_) return false;
        f.sin_t = tab_->sin_t.devPtr<float>(); f.cos_t = tab_->cos_t.devPtr<float>();
        f.rope_type = pose_type_; f.rot_dim = ishape_; f.pos = h_cnt_; f.pos_max = pos_max_;
        return true;
    }
    void advance(int tokens) override { h_cnt_ += tokens; }

private:
    void build(RopeTables &t) {
        const int dim = ishape_;
        vector<float> theta(dim / 2), s((size_t)pos_max_ * dim), c((size_t)pos_max_ * dim);
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
        for (int p = 0; p < pos_max_; p++) {
            for (int i = 0; i < dim / 2; i++) {
                const float ang = p * theta[i];
                const float sv = sinf(ang), cv = cosf(ang);
                const int d0 = pose_type_ == LLAMAROPE ? 2 * i : i;
                const int d1 = pose_type_ == LLAMAROPE ? 2 * i + 1 : i + dim / 2;
                s[(size_t)p * dim + d0] = sv; c[(size_t)p * dim + d0] = cv;
                s[(size_t)p * dim + d1] = sv; c[(size_t)p * dim + d1] = cv;
            }
        }
        t.dim = dim;
        t.pos_max = pos_max_;
        t.sin_t = Tensor(1, 1, pos_max_, dim, backend_);
        t.cos_t = Tensor(1, 1, pos_max_, dim, backend_);
        t.sin_t.copyFromHost(s.data());
        t.cos_t.copyFromHost(c.data());
    }
    int pose_type_, pos_max_, ishape_ = 0, h_cnt_ = 0;
    float rope_theta_, partial_;
    int theta_type_ = 0;
    float l3_factor_ = 8.f, l3_low_ = 1.f, l3_high_ = 4.f, l3_old_ctx_ = 8192.f;
    shared_ptr<RopeTables> tab_;
};

// ----------------------------------------------------------------- KVCache
// fp16 cache [B][cache_limit][H][D] (x86 KVCache_TYPE = 16, Types.hpp:26-31; n_rep forced to 1
// under flash_attention_2, CPUKVCache.cpp:45-47).  Producers (k_rope, v_proj) write their
// fp16 output straight into the slice at cache_seq_len_ through the alias registered at
// trace time; execute() then only advances cache_seq_len_ (CPUKVCache.cpp:137-247).
class CUDAKVCache final : public Op {
public:
    CUDAKVCache(Backend *bn, string name, int head, int hidden, int n_rep, bool fa2, int cache_max)
        : Op(bn, name), head_(head), hidden_(hidden), n_rep_(fa2 ? 1 : n_rep), limit_(cache_max) {
        if (!fa2 && n_rep > 1) throw std::runtime_error("[mllm-cuda] KVCache " + name + ": eager GQA head replication not supported; use flash_attention_2");
    }
    ErrorCode reshape(vector<shared_ptr<Tensor>> in, vector<shared_ptr<Tensor>> out) override {
        const int B = in[0]->batch(), H = in[0]->head(), S = in[0]->sequence(), D = in[0]->dimension();
        ensure(B, H, D);
        const int total = S + seq_;
        if (total > limit_) { // CPUKVCache.cpp:121-126 prints this and exit(1)s; a library throws instead
            std::cerr << "\n[ERROR]: Current tokens exceed cache limit: " << total << ">" << limit_ << ";"
                      << "\n         Please set args `--limits` >" << limit_ << std::endl;
            throw std::runtime_error("[mllm-cuda] KV cache limit exceeded");
        }
        out[0]->shallowCopyFrom(cache_, 0, total);
        return MLLM_NO_ERROR;
    }
    // two uses: (trace) register the producer's output name as an alias of the next slice;
    // (alias) called by runOp with inputs = {} to place a producer's output in the slice.
    ErrorCode setUp(vector<shared_ptr<Tensor>> in, vector<shared_ptr<Tensor>> out) override {
        if (!in.empty()) {
            cb(backend_)->cache_alias[in[0]->origin()] = this;
            return MLLM_NO_ERROR;
        }
        Tensor &o = *out[0];
        if (!cache_.allocated() || cache_.batch() != o.batch()) ensure(o.batch(), H_, D_);
        if (o.sequence() + seq_ > limit_) throw std::runtime_error("[mllm-cuda] KV cache limit exceeded");
        o.aliasInto(cache_, seq_);
        return MLLM_NO_ERROR;
    }
    ErrorCode execute(vector<shared_ptr<Tensor>> in, vector<shared_ptr<Tensor>> out) override {
        auto *b = cb(backend_);
        Tensor &x = *in[0];
        if (x.impl_ != cache_.impl_) { // not produced in place: convert + append
            if (x.dtype() != MLLM_TYPE_F32) throw std::runtime_error("[mllm-cuda] KVCache append expects fp32 input");
            for (int i = 0; i < x.batch(); i++) {
                __fp16_ptr dst = (char *)cache_.rawDevPtr() + ((size_t)i * cache_.batchStride() + (size_t)seq_ * cache_.seqStride()) * 2;
                MLLMCU_CHECK(mllmcu_f32_to_f16_rows(x.devPtr<float>() + (size_t)i * x.batchStride(), dst, x.sequence(), x.head() * x.dimension(),
                                                    x.seqStride(), cache_.seqStride(), b->stream()));
                b->kernel_launches++;
            }
        }
        seq_ += x.sequence();
        return MLLM_NO_ERROR;
    }
    int getCacheSeqLen() override { return seq_; }
    void clearCache() override { seq_ = 0; }
    bool fuseInfo(FuseInfo &f) override {
        if (!cache_.allocated()) return false;
        f.cache = cache_.rawDevPtr(); f.cache_bstride = cache_.batchStride(); f.cache_limit = limit_; f.cache_len = seq_;
        f.cache_batch = cache_.batch(); f.cache_heads = H_ * n_rep_; f.cache_dim = D_;
        return true;
    }
    void advance(int tokens) override { seq_ += tokens; }

private:
    typedef char *__fp16_ptr;
    void ensure(int B, int H, int D) {
        if (cache_.allocated() && cache_.batch() == B) return;
        H_ = H; D_ = D;
        cache_ = Tensor(backend_);
        cache_.setName(name() + ".Cache");
        cache_.setDtype(MLLM_TYPE_F16);
        cache_.reshape(B, H * n_rep_, limit_, D);
        cache_.alloc();
        MLLMCU_CHECK(mllmcu_memset(cache_.rawDevPtr(), 0, cache_.cntSize(), cb(backend_)->stream())); // CPUKVCache.cpp:53-66
        seq_ = 0;
    }
    int head_, hidden_, n_rep_, limit_, seq_ = 0, H_ = 0, D_ = 0;
    Tensor cache_;
};

// --------------------------------------------------------------- Embedding
class CUDAEmbedding final : public Op {
public:
    CUDAEmbedding(Backend *bn, string name, int hidden, int vocab) : Op(bn, name), hidden_(hidden), vocab_(vocab) {}
    ErrorCode reshape(vector<shared_ptr<Tensor>> in, vector<shared_ptr<Tensor>> out) override {
        out[0]->reshape(in[0]->batch(), 1, in[0]->sequence(), hidden_);
        return MLLM_NO_ERROR;
    }
    ErrorCode load(ParamLoader &loader) override {
        weight_ = Tensor(backend_);
        weight_.setName(name() + ".weight");
        weight_.reshape(1, 1, vocab_, hidden_);
        DataType dt = loader.getDataType(weight_.name());
        const bool absent = dt == MLLM_TYPE_COUNT;
        if (absent) dt = MLLM_TYPE_Q4_0;
        if (dt != MLLM_TYPE_Q4_0 && dt != MLLM_TYPE_F32) throw std::runtime_error("[mllm-cuda] Embedding " + name() + ": Q4_0 and F32 tables only");
        weight_.setDtype(dt);
        weight_.alloc();
        if (!absent) loader.load(&weight_);
        return MLLM_NO_ERROR;
    }
    ErrorCode execute(vector<shared_ptr<Tensor>> in, vector<shared_ptr<Tensor>> out) override {
        auto *b = cb(backend_);
        const int n = in[0]->batch() * in[0]->sequence();
        MLLMCU_CHECK(mllmcu_memset(out[0]->rawDevPtr(), 0, out[0]->cntSize(), b->stream()));
        if (weight_.dtype() == MLLM_TYPE_F32) MLLMCU_CHECK(mllmcu_embedding_f32(weight_.rawDevPtr(), in[0]->devPtr<float>(), out[0]->devPtr<float>(), n, hidden_, b->stream()));
        else MLLMCU_CHECK(mllmcu_embedding_q4_0(weight_.rawDevPtr(), in[0]->devPtr<float>(), out[0]->devPtr<float>(), n, hidden_, b->stream()));
        b->kernel_launches++;
        return MLLM_NO_ERROR;
    }

    bool fuseInfo(FuseInfo &f) override {
        if (!weight_.allocated()) return false;
        f.weight = weight_.rawDevPtr(); f.wtype = (int)weight_.dtype(); f.in = vocab_; f.out = hidden_;
        return true;
    }

private:
    int hidden_, vocab_;
    Tensor weight_;
};

// ------------------------------------------------- SiLU / binary functions
class CUDASiLU final : public Op {
public:
    CUDASiLU(Backend *bn, string name) : Op(bn, name) {}
    ErrorCode reshape(vector<shared_ptr<Tensor>> in, vector<shared_ptr<Tensor>> out) override {
        out[0]->reshape(in[0]->batch(), in[0]->head(), in[0]->sequence(), in[0]->dimension());
        return MLLM_NO_ERROR;
    }
    ErrorCode execute(vector<shared_ptr<Tensor>> in, vector<shared_ptr<Tensor>> out) override {
        auto *b = cb(backend_);
        MLLMCU_CHECK(mllmcu_silu(in[0]->devPtr<float>(), out[0]->devPtr<float>(), in[0]->count(), b->stream()));
        b->kernel_launches++;
        return MLLM_NO_ERROR;
    }
};

class CUDABinary final : public Op { // F_TTADD / F_TTMUL
public:
    CUDABinary(Backend *bn, string name, bool is_mul) : Op(bn, name), mul_(is_mul) {}
    ErrorCode reshape(vector<shared_ptr<Tensor>> in, vector<shared_ptr<Tensor>> out) override {
        if (in[0]->count() != in[1]->count()) throw std::runtime_error("[mllm-cuda] binary op: shape mismatch (broadcast not supported)");
        out[0]->reshape(in[0]->batch(), in[0]->head(), in[0]->sequence(), in[0]->dimension());
        return MLLM_NO_ERROR;
    }
    ErrorCode execute(vector<shared_ptr<Tensor>> in, vector<shared_ptr<Tensor>> out) override {
        auto *b = cb(backend_);
        if (mul_) MLLMCU_CHECK(mllmcu_mul(in[0]->devPtr<float>(), in[1]->devPtr<float>(), out[0]->devPtr<float>(), in[0]->count(), b->stream()));
        else MLLMCU_CHECK(mllmcu_add(in[0]->devPtr<float>(), in[1]->devPtr<float>(), out[0]->devPtr<float>(), in[0]->count(), b->stream()));
        b->kernel_launches++;
        return MLLM_NO_ERROR;
    }

private:
    bool mul_;
};

// -------------------------------------------------------------------- view
// Tensor::view(b,h,s,d) (Tensor.cpp F_VIEW): -1 keeps/infers.  In BSHD memory
// [B,1,S,H*D] and [B,H,S,D] are the same bytes, so this is metadata only.
class CUDAView final : public Op {
public:
    CUDAView(Backend *bn, string name, int b, int h, int s, int d) : Op(bn, name), vb_(b), vh_(h), vs_(s), vd_(d) {}
    ErrorCode reshape(vector<shared_ptr<Tensor>> in, vector<shared_ptr<Tensor>> out) override {
        Tensor &x = *in[0];
        const int B = vb_ == -1 ? x.batch() : vb_;
        const int S = vs_ == -1 ? x.sequence() : vs_;
        int H = vh_, D = vd_;
        const long long hd = (long long)x.head() * x.dimension();
        if (H == -1 && D == -1) { H = x.head(); D = x.dimension(); }
        else if (H == -1) H = (int)(hd / D);
        else if (D == -1) D = (int)(hd / H);
        if ((long long)H * D != hd || B != x.batch() || S != x.sequence()) throw std::runtime_error("[mllm-cuda] view: only head/dimension re-splits are supported");
        const string nm = out[0]->name();
        *out[0] = x; // shares storage (and strides: a cache-slice alias stays one)
        out[0]->setName(nm);
        out[0]->setOrigin(x.origin());
        out[0]->reshapeKeepStrides(B, H, S, D);
        return MLLM_NO_ERROR;
    }

private:
    int vb_, vh_, vs_, vd_;
};

// -------------------------------------------------------------------- clip
// x.clip({}, {}, {-1}, {}): keep the last sequence position (modeling_llama.hpp:109-112).
class CUDAClip final : public Op {
public:
    CUDAClip(Backend *bn, string name, int s_index) : Op(bn, name), s_(s_index) {}
    ErrorCode reshape(vector<shared_ptr<Tensor>> in, vector<shared_ptr<Tensor>> out) override {
        out[0]->reshape(in[0]->batch(), in[0]->head(), 1, in[0]->dimension());
        return MLLM_NO_ERROR;
    }
    ErrorCode execute(vector<shared_ptr<Tensor>> in, vector<shared_ptr<Tensor>> out) override {
        auto *b = cb(backend_);
        Tensor &x = *in[0];
        const int s = s_ < 0 ? x.sequence() + s_ : s_;
        const size_t row = (size_t)x.head() * x.dimension() * 4;
        for (int i = 0; i < x.batch(); i++)
            MLLMCU_CHECK(mllmcu_memcpy_d2d((char *)out[0]->rawDevPtr() + i * row,
                                           (char *)x.rawDevPtr() + ((size_t)i * x.batchStride() + (size_t)s * x.seqStride()) * 4, row, b->stream()));
        return MLLM_NO_ERROR;
    }

private:
    int s_;
};

// ---------------------------------------------------------- flash attention
class CUDAFlashAttention2 final : public Op {
public:
    CUDAFlashAttention2(Backend *bn, string name, bool causal) : Op(bn, name), causal_(causal) {}
    ErrorCode reshape(vector<shared_ptr<Tensor>> in, vector<shared_ptr<Tensor>> out) override {
        out[0]->reshape(in[0]->batch(), in[0]->head(), in[0]->sequence(), in[0]->dimension());
        return MLLM_NO_ERROR;
    }
    ErrorCode execute(vector<shared_ptr<Tensor>> in, vector<shared_ptr<Tensor>> out) override {
        auto *b = cb(backend_);
        Tensor &q = *in[0], &k = *in[1], &v = *in[2];
        if (k.dtype() != v.dtype() || (k.dtype() != MLLM_TYPE_F16 && k.dtype() != MLLM_TYPE_F32))
            throw std::runtime_error("[mllm-cuda] flash_attention2: K/V must both be fp16 (KV cache) or both fp32 (no cache: vision towers)");
        if (k.seqStride() != (long long)k.head() * k.dimension()) throw std::runtime_error("[mllm-cuda] flash_attention2: K/V rows must be dense");
        const size_t wsb = k.dtype() == MLLM_TYPE_F32 ? mllmcu_flash_attn_f32kv_workspace_bytes(q.batch(), q.sequence(), k.sequence(), q.head(), k.head(), q.dimension())
                                                      : mllmcu_flash_attn_workspace_bytes(q.batch(), q.sequence(), q.head(), q.dimension());
        void *ws = b->workspace(wsb);
        if (k.dtype() == MLLM_TYPE_F32) { // flash_attention_2_forward's fp32 variant (compute/FlashAttention2.hpp:87-1210): ViT blocks keep no cache
            MLLMCU_CHECK(mllmcu_flash_attn_f32kv(q.devPtr<float>(), k.devPtr<float>(), v.devPtr<float>(), out[0]->devPtr<float>(), q.batch(), q.sequence(),
                                                 k.sequence(), q.head(), k.head(), q.dimension(), causal_, k.batchStride(), ws, b->workspaceBytes(), b->stream()));
            b->kernel_launches += 1;
            return MLLM_NO_ERROR;
        }
        MLLMCU_CHECK(mllmcu_flash_attn_f16kv(q.devPtr<float>(), k.rawDevPtr(), v.rawDevPtr(), out[0]->devPtr<float>(), q.batch(), q.sequence(), k.sequence(),
                                             q.head(), k.head(), q.dimension(), causal_, k.batchStride(), ws, b->workspaceBytes(), b->stream()));
        b->kernel_launches += 2;
        return MLLM_NO_ERROR;
    }

private:
    bool causal_;
};


// ------------------------------------------------------ GELU / QuickGELU
// host-side fp16 <-> fp32 (IEEE, round-to-nearest-even) for building the lookup tables exactly as init_table_gelu_f16 does
static float h2f(uint16_t h) {
    const uint32_t sign = (uint32_t)(h & 0x8000) << 16;
    uint32_t e = (h >> 10) & 0x1F, m = h & 0x3FF, u;
    if (e == 0) {
        if (m == 0) u = sign;
        else { int sh = 0; while (!(m & 0x400)) { m <<= 1; sh++; } u = sign | ((uint32_t)(113 - sh) << 23) | ((m & 0x3FF) << 13); }
    } else if (e == 31) u = sign | 0x7F800000u | (m << 13) | (m ? 0x400000u : 0);
    else u = sign | ((e + 112) << 23) | (m << 13);
    float f; memcpy(&f, &u, 4); return f;
}
static uint16_t f2h(float f) {
    uint32_t u; memcpy(&u, &f, 4);
    const uint16_t sign = (uint16_t)((u >> 16) & 0x8000);
    u &= 0x7FFFFFFFu;
    if (u >= 0x7F800000u) return sign | 0x7C00 | (u > 0x7F800000u ? (0x200 | ((u >> 13) & 0x3FF)) : 0);
    if (u >= 0x477FF000u) return sign | 0x7C00; // rounds to >= 65520 -> inf
    if (u < 0x33000001u) return sign;           // < 2^-25 (or exactly 2^-25, ties to even 0)
    const int e = (int)(u >> 23) - 127;
    uint32_t m = (u & 0x7FFFFFu) | 0x800000u;
    int shift = e < -14 ? (13 + (-14 - e)) : 13;
    uint32_t r = m >> shift, rem = m & ((1u << shift) - 1), halfv = 1u << (shift - 1);
    if (rem > halfv || (rem == halfv && (r & 1))) r++;
    if (e < -14) return sign | (uint16_t)r; // subnormal (r may carry into the smallest normal: same bits)
    return sign | (uint16_t)(((uint32_t)(e + 15) << 10) + (r - 0x400));
}

class CUDAGelu final : public Op { // kind 0: mllm_vec_gelu_f32, 1: mllm_vec_gelu_quick_f32 (ggml/Quantize.hpp:74-133)
public:
    CUDAGelu(Backend *bn, string name, int kind) : Op(bn, name), kind_(kind) {}
    ErrorCode reshape(vector<shared_ptr<Tensor>> in, vector<shared_ptr<Tensor>> out) override {
        out[0]->reshape(in[0]->batch(), in[0]->head(), in[0]->sequence(), in[0]->dimension());
        return MLLM_NO_ERROR;
    }
    ErrorCode execute(vector<shared_ptr<Tensor>> in, vector<shared_ptr<Tensor>> out) override {
        auto *b = cb(backend_);
        if (!table_.allocated()) { // built once per op, like the reference's lazily initialised static tables
            std::vector<uint16_t> t(1 << 16);
            for (int i = 0; i < (1 << 16); i++) {
                const float x = h2f((uint16_t)i);
                // gelu: g++ -O3 with FMA contracts 1 + (A*x)*x into one fma in the reference build (pinned by tests/golden)
                const float g = kind_ == 0 ? 0.5f * x * (1.0f + tanhf(0.79788456080286535587989211986876f * x * std::fmaf(0.044715f * x, x, 1.0f)))
                                           : x * (1.0f / (1.0f + expf(-1.702f * x)));
                t[i] = f2h(g);
            }
            table_ = Tensor(1, 1, 1, 1 << 16, backend_, false);
            table_.setDtype(MLLM_TYPE_F16);
            table_.alloc();
            table_.copyFromHost(t.data());
        }
        MLLMCU_CHECK(mllmcu_gelu_table(in[0]->devPtr<float>(), out[0]->devPtr<float>(), in[0]->count(), table_.rawDevPtr(), b->stream()));
        b->kernel_launches++;
        return MLLM_NO_ERROR;
    }

private:
    int kind_;
    Tensor table_;
};

// ---------------------------------------------------------------- SoftMax
// CPUSoftMax (op/CPUSoftMax.cpp:29-66), axis = DIMENSION.  In this backend's BSHD memory a [B,H,S,D] score tensor has its
// rows at (b, s, h): the kernel's row index only needs s for the causal length, so rows are visited as outer = B, S, n = D per
// head by running one launch per head slice when H > 1 would break contiguity -- scores produced by CUDAMatmulFunc are stored
// head-major ([B][H][S][D] dense, flagged through origin "bhsd"), which is what this op expects.
class CUDASoftMax final : public Op {
public:
    CUDASoftMax(Backend *bn, string name, int axis, bool causal) : Op(bn, name), axis_(axis), causal_(causal) {}
    ErrorCode reshape(vector<shared_ptr<Tensor>> in, vector<shared_ptr<Tensor>> out) override {
        out[0]->reshape(in[0]->batch(), in[0]->head(), in[0]->sequence(), in[0]->dimension());
        return MLLM_NO_ERROR;
    }
    ErrorCode execute(vector<shared_ptr<Tensor>> in, vector<shared_ptr<Tensor>> out) override {
        auto *b = cb(backend_);
        Tensor &x = *in[0];
        if (axis_ != 3) throw std::runtime_error("[mllm-cuda] softmax: only axis = DIMENSION");
        if (x.head() != 1) throw std::runtime_error("[mllm-cuda] softmax: scores must be [B*H,1,S,D] head-major rows (see CUDAMatmulFunc)");
        const int old_dim = x.dimension() - x.sequence(); // CPUSoftMax.cpp:38-40
        MLLMCU_CHECK(mllmcu_softmax(x.devPtr<float>(), out[0]->devPtr<float>(), x.batch(), x.sequence(), x.dimension(), causal_, old_dim < 0 ? 0 : old_dim, b->stream()));
        b->kernel_launches++;
        return MLLM_NO_ERROR;
    }

private:
    int axis_;
    bool causal_;
};

// x / scalar: Tensor::operator/(float) -> F_DIV -> CPUdivFunction (Tensor.cpp:445-450, op/CPUBinaryFunc.hpp)
class CUDADivScalar final : public Op {
public:
    CUDADivScalar(Backend *bn, string name, float v) : Op(bn, name), v_(v) {}
    ErrorCode reshape(vector<shared_ptr<Tensor>> in, vector<shared_ptr<Tensor>> out) override {
        out[0]->reshape(in[0]->batch(), in[0]->head(), in[0]->sequence(), in[0]->dimension());
        return MLLM_NO_ERROR;
    }
    ErrorCode execute(vector<shared_ptr<Tensor>> in, vector<shared_ptr<Tensor>> out) override {
        auto *b = cb(backend_);
        MLLMCU_CHECK(mllmcu_div_scalar(in[0]->devPtr<float>(), v_, out[0]->devPtr<float>(), in[0]->count(), b->stream()));
        b->kernel_launches++;
        return MLLM_NO_ERROR;
    }

private:
    float v_;
};

// Tensor::mm (F_MM, op/CPUMatmulFunc.hpp) on fp32 operands laid out head-major: a [G,1,M,K] x b [G,1,N,K]^T (transpose1) or
// b [G,1,K,N] -> [G,1,M,N], G = batch*heads.  The eager attention spelling reaches it after CUDATranspose (BSHD -> BHSD).
class CUDAMatmulFunc final : public Op {
public:
    CUDAMatmulFunc(Backend *bn, string name, bool t1) : Op(bn, name), t1_(t1) {}
    ErrorCode reshape(vector<shared_ptr<Tensor>> in, vector<shared_ptr<Tensor>> out) override {
        Tensor &a = *in[0], &bm = *in[1];
        if (a.head() != 1 || bm.head() != 1 || a.batch() != bm.batch()) throw std::runtime_error("[mllm-cuda] mm: operands must be [G,1,rows,cols]");
        const int N = t1_ ? bm.sequence() : bm.dimension();
        if ((t1_ ? bm.dimension() : bm.sequence()) != a.dimension()) throw std::runtime_error("[mllm-cuda] mm: inner dimensions differ");
        out[0]->reshape(a.batch(), 1, a.sequence(), N);
        return MLLM_NO_ERROR;
    }
    ErrorCode execute(vector<shared_ptr<Tensor>> in, vector<shared_ptr<Tensor>> out) override {
        auto *b = cb(backend_);
        Tensor &a = *in[0], &bm = *in[1], &c = *out[0];
        const int M = a.sequence(), K = a.dimension(), N = c.dimension();
        MLLMCU_CHECK(mllmcu_sgemm(a.devPtr<float>(), a.seqStride(), a.batchStride(), bm.devPtr<float>(), t1_ ? bm.seqStride() : 1, t1_ ? 1 : bm.seqStride(),
                                  bm.batchStride(), nullptr, c.devPtr<float>(), c.seqStride(), c.batchStride(), M, N, K, a.batch(), b->stream()));
        b->kernel_launches++;
        return MLLM_NO_ERROR;
    }

private:
    bool t1_;
};

// x.transpose(HEAD, SEQUENCE) as a physical copy (op/CPUTransposeFunc.hpp:127-149): BSHD [B,H,S,D] -> head-major rows,
// returned as [B*H, 1, S, D] so the matmul / softmax ops above see dense per-head matrices.  Works on cache-slice views.
class CUDATranspose final : public Op {
public:
    CUDATranspose(Backend *bn, string name) : Op(bn, name) {}
    ErrorCode reshape(vector<shared_ptr<Tensor>> in, vector<shared_ptr<Tensor>> out) override {
        out[0]->reshape(in[0]->batch() * in[0]->head(), 1, in[0]->sequence(), in[0]->dimension());
        out[0]->setDtype(in[0]->dtype());
        return MLLM_NO_ERROR;
    }
    ErrorCode execute(vector<shared_ptr<Tensor>> in, vector<shared_ptr<Tensor>> out) override {
        auto *b = cb(backend_);
        Tensor &x = *in[0];
        const int n[4] = {x.batch(), x.head(), x.sequence(), x.dimension()};
        const long long ss[4] = {x.batchStride(), x.dimension(), x.seqStride(), 1};
        const long long ds[4] = {(long long)x.head() * x.sequence() * x.dimension(), (long long)x.sequence() * x.dimension(), x.dimension(), 1};
        MLLMCU_CHECK(mllmcu_copy_strided(x.rawDevPtr(), out[0]->rawDevPtr(), x.dtype() == MLLM_TYPE_F16 ? 2 : 4, n, ss, ds, b->stream()));
        b->kernel_launches++;
        return MLLM_NO_ERROR;
    }
};

// ---------------------------------------------------------------- Parameter
// CPUParameter (op/CPUParameter.cpp:8-60): a learned tensor loaded under the op's own name (ViT cls_token and
// position_embeddings, models/vit/modeling_vit.hpp:66-84); no inputs, the output is the weight.
class CUDAParameter final : public Op {
public:
    CUDAParameter(Backend *bn, string name, int b, int h, int s, int d) : Op(bn, name), b_(b), h_(h), s_(s), d_(d) {}
    ErrorCode reshape(vector<shared_ptr<Tensor>>, vector<shared_ptr<Tensor>> out) override {
        out[0]->reshape(b_, h_, s_, d_);
        return MLLM_NO_ERROR;
    }
    ErrorCode load(ParamLoader &loader) override {
        w_ = Tensor(backend_);
        w_.setName(name());
        w_.reshape(b_, h_, s_, d_);
        w_.setDtype(MLLM_TYPE_F32);
        w_.alloc();
        if (loader.getDataType(w_.name()) != MLLM_TYPE_COUNT) loader.load(&w_);
        return MLLM_NO_ERROR;
    }
    ErrorCode execute(vector<shared_ptr<Tensor>>, vector<shared_ptr<Tensor>> out) override {
        auto *b = cb(backend_);
        MLLMCU_CHECK(mllmcu_memcpy_d2d(out[0]->rawDevPtr(), w_.rawDevPtr(), w_.cntSize(), b->stream()));
        return MLLM_NO_ERROR;
    }

private:
    int b_, h_, s_, d_;
    Tensor w_;
};

// ---------------------------------------------------------------------- Cat
// CPUcatFunction (op/CPUCatFunc.hpp:14-260): concatenate along BATCH / SEQUENCE / HEAD / DIMENSION (Chl values 0 / 1 / 2 / 3,
// Types.hpp:133-137).  The reference aliases the inputs into the output where it can; the result is the same bytes.  Here
// every input is placed with one strided copy (cls token + patches, image + text embeddings, rotary halves).
class CUDACat final : public Op {
public:
    CUDACat(Backend *bn, string name, int axis) : Op(bn, name), axis_(axis) {
        if (axis_ < 0 || axis_ > 3) throw std::runtime_error("[mllm-cuda] cat: axis must be BATCH, SEQUENCE, HEAD or DIMENSION");
    }
    ErrorCode reshape(vector<shared_ptr<Tensor>> in, vector<shared_ptr<Tensor>> out) override {
        int n[4] = {in[0]->batch(), in[0]->sequence(), in[0]->head(), in[0]->dimension()}; // indexed by Chl
        n[axis_] = 0;
        for (auto &t : in) {
            const int v[4] = {t->batch(), t->sequence(), t->head(), t->dimension()};
            for (int a = 0; a < 4; a++)
                if (a != axis_ && v[a] != n[a]) throw std::runtime_error("[mllm-cuda] cat: shapes differ off the concatenation axis");
            n[axis_] += v[axis_];
        }
        out[0]->reshape(n[0], n[2], n[1], n[3]);
        out[0]->setDtype(in[0]->dtype());
        return MLLM_NO_ERROR;
    }
    ErrorCode execute(vector<shared_ptr<Tensor>> in, vector<shared_ptr<Tensor>> out) override {
        auto *b = cb(backend_);
        Tensor &y = *out[0];
        const int esz = y.dtype() == MLLM_TYPE_F16 ? 2 : 4;
        // element (b, s, h, d) of a BSHD tensor: b * batchStride + s * seqStride + h * D + d
        const long long ys[4] = {y.batchStride(), y.seqStride(), y.dimension(), 1};
        long long at = 0;
        for (auto &t : in) {
            const int n[4] = {t->batch(), t->sequence(), t->head(), t->dimension()};
            const long long ss[4] = {t->batchStride(), t->seqStride(), t->dimension(), 1};
            MLLMCU_CHECK(mllmcu_copy_strided(t->rawDevPtr(), (char *)y.rawDevPtr() + (size_t)(at * ys[axis_]) * esz, esz, n, ss, ys, b->stream()));
            b->kernel_launches++;
            at += n[axis_];
        }
        return MLLM_NO_ERROR;
    }

private:
    int axis_;
};

// -------------------------------------------------------------------- Split
// CPUsplitFunction (op/CPUSplitFunc.hpp:18-175): cut the input into pieces of each_dims along HEAD / SEQUENCE / DIMENSION (Chl 2 / 1 / 3).
// The merged-qkv layouts D_HD / HD (Chl 313 / 113) are not built.
class CUDASplit final : public Op {
public:
    CUDASplit(Backend *bn, string name, vector<int> each, int axis) : Op(bn, name), each_(std::move(each)), axis_(axis) {
        if (axis_ < 1 || axis_ > 3) throw std::runtime_error("[mllm-cuda] split: HEAD, SEQUENCE or DIMENSION only");
    }
    ErrorCode reshape(vector<shared_ptr<Tensor>> in, vector<shared_ptr<Tensor>> out) override {
        if (out.size() != each_.size()) throw std::runtime_error("[mllm-cuda] split: one output per piece");
        const int n[4] = {in[0]->batch(), in[0]->sequence(), in[0]->head(), in[0]->dimension()}; // indexed by Chl
        int total = 0;
        for (int e : each_) total += e;
        if (total != n[axis_]) throw std::runtime_error("[mllm-cuda] split: pieces do not add up to the axis length");
        for (size_t i = 0; i < out.size(); i++) {
            int m[4] = {n[0], n[1], n[2], n[3]};
            m[axis_] = each_[i];
            out[i]->reshape(m[0], m[2], m[1], m[3]);
            out[i]->setDtype(in[0]->dtype());
        }
        return MLLM_NO_ERROR;
    }
    ErrorCode execute(vector<shared_ptr<Tensor>> in, vector<shared_ptr<Tensor>> out) override {
        auto *b = cb(backend_);
        Tensor &x = *in[0];
        const int esz = x.dtype() == MLLM_TYPE_F16 ? 2 : 4;
        const long long xs[4] = {x.batchStride(), x.seqStride(), x.dimension(), 1};
        long long at = 0;
        for (size_t i = 0; i < out.size(); i++) {
            Tensor &y = *out[i];
            const int n[4] = {y.batch(), y.sequence(), y.head(), y.dimension()};
            const long long ys[4] = {y.batchStride(), y.seqStride(), y.dimension(), 1};
            MLLMCU_CHECK(mllmcu_copy_strided((const char *)x.rawDevPtr() + (size_t)(at * xs[axis_]) * esz, y.rawDevPtr(), esz, n, xs, ys, b->stream()));
            b->kernel_launches++;
            at += n[axis_];
        }
        return MLLM_NO_ERROR;
    }

private:
    vector<int> each_;
    int axis_;
};
static Op *make_split(const OpParam &p, Backend *bn, const string &name) {
    vector<int> each;
    for (int i = 0; i < (int)p.at("num_splits"); i++) each.push_back((int)p.at("dim_" + std::to_string(i)));
    return new CUDASplit(bn, name, each, (int)p.at("split_dim"));
}

// ------------------------------------------------------------ Convolution2D
// CPUConvolution2D (op/CPUConvolution2D.cpp) VALID padding: input [B, H=height, S=channel, D=width] (NCHW bytes in BSHD), weight
// name + ".weight" [O][C][kh][kw] F32, optional bias; output [B, H=oh, S=O, D=ow].
class CUDAConvolution2D final : public Op {
public:
    CUDAConvolution2D(Backend *bn, string name, const OpParam &p)
        : Op(bn, name), C_((int)p.at("in_channel")), O_((int)p.at("out_channel")), kh_((int)p.at("kernal_h")), kw_((int)p.at("kernal_w")),
          sh_((int)p.at("stride_h")), sw_((int)p.at("stride_w")), pad_((int)getp(p, "padding", 1)), bias_((bool)getp(p, "bias", 1)) {
        if (pad_ != 1) throw std::runtime_error("[mllm-cuda] Convolution2D: only VALID padding (the patch-embed case)"); // PaddingType VALID = 1 (Types.hpp)
    }
    ErrorCode reshape(vector<shared_ptr<Tensor>> in, vector<shared_ptr<Tensor>> out) override {
        if (in[0]->sequence() != C_) throw std::runtime_error("[mllm-cuda] Convolution2D: channel mismatch");
        out[0]->reshape(in[0]->batch(), (in[0]->head() - kh_) / sh_ + 1, O_, (in[0]->dimension() - kw_) / sw_ + 1);
        return MLLM_NO_ERROR;
    }
    ErrorCode load(ParamLoader &loader) override {
        w_ = Tensor(backend_);
        w_.setName(name() + ".weight");
        w_.reshape(O_, kh_, C_, kw_);
        w_.setDtype(MLLM_TYPE_F32);
        w_.alloc();
        if (loader.getDataType(w_.name()) != MLLM_TYPE_COUNT) loader.load(&w_);
        if (bias_) {
            b_ = Tensor(backend_);
            b_.setName(name() + ".bias");
            b_.reshape(1, 1, 1, O_);
            b_.setDtype(MLLM_TYPE_F32);
            b_.alloc();
            if (loader.getDataType(b_.name()) != MLLM_TYPE_COUNT) loader.load(&b_);
        }
        return MLLM_NO_ERROR;
    }
    ErrorCode execute(vector<shared_ptr<Tensor>> in, vector<shared_ptr<Tensor>> out) override {
        auto *b = cb(backend_);
        Tensor &x = *in[0];
        MLLMCU_CHECK(mllmcu_conv2d_valid(x.devPtr<float>(), w_.devPtr<float>(), bias_ ? b_.devPtr<float>() : nullptr, out[0]->devPtr<float>(), x.batch(), C_,
                                         x.head(), x.dimension(), O_, kh_, kw_, sh_, sw_, 0, b->stream()));
        b->kernel_launches++;
        return MLLM_NO_ERROR;
    }

private:
    int C_, O_, kh_, kw_, sh_, sw_, pad_;
    bool bias_;
    Tensor w_, b_;
};

// ----------------------------------------------------------- MultimodalRoPE
// CPUMultimodalRoPE (op/CPUMultimodalRoPE.cpp:25-331): inputs {x [B,H,S,D], position_ids [3,1,1,S]}.  The sin/cos tables are
// rebuilt on the HOST from the ids exactly as multimodal_sinusoidal_position_embedding + apply_multimodal_rotary_pos_emb do
// (same libm sinf/cosf, so the rotation is bit-identical to the reference), uploaded, and applied rotate-half by row s.
class CUDAMultimodalRoPE final : public Op {
public:
    CUDAMultimodalRoPE(Backend *bn, string name, const OpParam &p) : Op(bn, name), theta_((float)p.at("rope_theta")) {
        for (int i = 0;; i++) {
            auto it = p.find("mrope_section_" + std::to_string(i));
            if (it == p.end()) break;
            sec_.push_back((int)it->second);
        }
        // {16,24,24} (configuration_qwen2_vl.hpp:48) covers the D/2 frequency columns the rotation reads; the rest of the table stays 0
    }
    ErrorCode reshape(vector<shared_ptr<Tensor>> in, vector<shared_ptr<Tensor>> out) override {
        out[0]->reshape(in[0]->batch(), in[0]->head(), in[0]->sequence(), in[0]->dimension());
        return MLLM_NO_ERROR;
    }
    ErrorCode execute(vector<shared_ptr<Tensor>> in, vector<shared_ptr<Tensor>> out) override {
        auto *b = cb(backend_);
        Tensor &x = *in[0], &pid = *in[1];
        const int S = x.sequence(), D = x.dimension(), half = D / 2;
        if (pid.batch() != 3 || pid.dimension() < S) throw std::runtime_error("[mllm-cuda] MultimodalRoPE: position_ids must be [3,1,1,S]");
        std::vector<float> pos((size_t)3 * pid.dimension());
        b->sync();
        pid.copyToHost(pos.data());
        std::vector<float> theta(half), sn((size_t)S * D, 0.f), cs((size_t)S * D, 0.f);
        for (int i = 0; i < half; i++) theta[i] = (float)(1.0 / pow((double)theta_, 2.0 * i / D));
        int start = 0;
        for (size_t j = 0; j < sec_.size(); j++) {
            const int axis = (int)(j % 3);
            for (int c = start; c < start + sec_[j] && c < D; c++)
                for (int s = 0; s < S; s++) {
                    const float value = theta[c % half] * pos[(size_t)axis * pid.dimension() + s];
                    cs[(size_t)s * D + c] = cosf(value) * 1.0f;
                    sn[(size_t)s * D + c] = sinf(value) * 1.0f;
                }
            start += sec_[j];
        }
        if (!sin_.allocated() || sin_.sequence() < S) {
            sin_ = Tensor(1, 1, S, D, backend_);
            cos_ = Tensor(1, 1, S, D, backend_);
        }
        MLLMCU_CHECK(mllmcu_memcpy_h2d(sin_.rawDevPtr(), sn.data(), sn.size() * 4, b->stream()));
        MLLMCU_CHECK(mllmcu_memcpy_h2d(cos_.rawDevPtr(), cs.data(), cs.size() * 4, b->stream()));
        b->sync(); // the host vectors die at scope exit
        Tensor &y = *out[0];
        const int f16 = y.dtype() == MLLM_TYPE_F16;
        MLLMCU_CHECK(mllmcu_rope_mrope(x.devPtr<float>(), y.rawDevPtr(), x.batch(), S, x.head(), D, D, sin_.devPtr<float>(), cos_.devPtr<float>(), D, f16,
                                       y.batchStride(), y.seqStride(), b->stream()));
        b->kernel_launches++;
        return MLLM_NO_ERROR;
    }

private:
    float theta_;
    vector<int> sec_;
    Tensor sin_, cos_;
};

// Tensor::apply_rotary_pos_emb_vision (F_APPLY_VISIOROPE, op/CPUVisionRoPEFunc.hpp:20-160): inputs {x [B,H,S,D], rotary_pos_emb
// [1,1,S,D/2]}; sin/cos of the angles are taken on the device (<= 2 ulp from libm; mllmcu_rope_vision also accepts host tables).
class CUDAVisionRoPEFunc final : public Op {
public:
    CUDAVisionRoPEFunc(Backend *bn, string name) : Op(bn, name) {}
    ErrorCode reshape(vector<shared_ptr<Tensor>> in, vector<shared_ptr<Tensor>> out) override {
        out[0]->reshape(in[0]->batch(), in[0]->head(), in[0]->sequence(), in[0]->dimension());
        return MLLM_NO_ERROR;
    }
    ErrorCode execute(vector<shared_ptr<Tensor>> in, vector<shared_ptr<Tensor>> out) override {
        auto *b = cb(backend_);
        Tensor &x = *in[0];
        if (in[1]->dimension() * 2 != x.dimension() || in[1]->sequence() != x.sequence()) throw std::runtime_error("[mllm-cuda] vision rope: rotary_pos_emb must be [1,1,S,D/2]");
        MLLMCU_CHECK(mllmcu_rope_vision(x.devPtr<float>(), in[1]->devPtr<float>(), nullptr, out[0]->devPtr<float>(), x.batch(), x.sequence(), x.head(), x.dimension(), 0, b->stream()));
        b->kernel_launches++;
        return MLLM_NO_ERROR;
    }
};

// ---------------------------------------------------------------- creators
#define CREATOR(cls, expr)                                                                    \
    class cls##Creator : public CUDABackend::Creator {                                        \
    public:                                                                                   \
        Op *create(OpParam p, Backend *bn, string name, int) const override { return expr; } \
    };
CREATOR(CUDALinear, new CUDALinear(bn, name, (int)p.at("in_features"), (int)p.at("out_features"), (bool)p.at("bias")))
CREATOR(CUDARMSNorm, new CUDARMSNorm(bn, name, (int)p.at("norm_size"), getp(p, "epsilon", 1e-6f), (bool)getp(p, "add_unit_offset", 0)))
CREATOR(CUDALayerNorm, new CUDALayerNorm(bn, name, (int)p.at("norm_size"), (bool)getp(p, "bias", 1), getp(p, "epsilon", 1e-6f)))
CREATOR(CUDARoPE, new CUDARoPE(bn, name, p))
CREATOR(CUDAKVCache, new CUDAKVCache(bn, name, (int)getp(p, "head", 0), (int)getp(p, "hidden", 0), (int)getp(p, "n_rep", 1), (bool)getp(p, "fa2", 0), (int)p.at("cache_max")))
CREATOR(CUDAEmbedding, new CUDAEmbedding(bn, name, (int)p.at("hidden_size"), (int)p.at("vocab_size")))
CREATOR(CUDASiLU, new CUDASiLU(bn, name))
CREATOR(CUDAAdd, new CUDABinary(bn, name, false))
CREATOR(CUDAMul, new CUDABinary(bn, name, true))
CREATOR(CUDAView, new CUDAView(bn, name, (int)getp(p, "b", -1), (int)getp(p, "h", -1), (int)getp(p, "s", -1), (int)getp(p, "d", -1)))
static Op *make_clip(const OpParam &p, Backend *bn, const string &name) {
    // Tensor::clip(b,h,s,d) packs its index lists as {b,h,s,d}_size + {x}_i keys (Tensor.cpp:533-551)
    if (getp(p, "b_size", 0) != 0 || getp(p, "h_size", 0) != 0 || getp(p, "d_size", 0) != 0 || getp(p, "s_size", 0) != 1)
        throw std::runtime_error("[mllm-cuda] clip: only a single sequence index is supported");
    return new CUDAClip(bn, name, (int)getp(p, "s_0", -1));
}
CREATOR(CUDAClip, make_clip(p, bn, name))
CREATOR(CUDAFlashAttention2, new CUDAFlashAttention2(bn, name, (bool)getp(p, "causal_mask", 1)))
CREATOR(CUDAGelu, new CUDAGelu(bn, name, 0))
CREATOR(CUDAQuickGelu, new CUDAGelu(bn, name, 1))
CREATOR(CUDASoftMax, new CUDASoftMax(bn, name, (int)getp(p, "axis", 3), (bool)getp(p, "do_causal_mask", 0)))
CREATOR(CUDADivScalar, new CUDADivScalar(bn, name, getp(p, "data", 1.f)))
CREATOR(CUDAMatmulFunc, new CUDAMatmulFunc(bn, name, (bool)getp(p, "transpose1", 0)))
CREATOR(CUDATranspose, new CUDATranspose(bn, name))
CREATOR(CUDAParameter, new CUDAParameter(bn, name, (int)p.at("batch"), (int)p.at("head"), (int)p.at("seq"), (int)p.at("dim")))
CREATOR(CUDACat, new CUDACat(bn, name, (int)p.at("axis")))
CREATOR(CUDASplit, make_split(p, bn, name))
CREATOR(CUDAConvolution2D, new CUDAConvolution2D(bn, name, p))
CREATOR(CUDAMultimodalRoPE, new CUDAMultimodalRoPE(bn, name, p))
CREATOR(CUDAVisionRoPEFunc, new CUDAVisionRoPEFunc(bn, name))

void CUDABackend::registerOps() { // CPUBackend::registerOps, CPUBackend.cpp:142-244
    addCreator(LINEAR, new CUDALinearCreator());
    addCreator(RMSNORM, new CUDARMSNormCreator());
    addCreator(LAYERNORM, new CUDALayerNormCreator());
    addCreator(ROPE, new CUDARoPECreator());
    addCreator(KVCACHE, new CUDAKVCacheCreator());
    addCreator(EMBEDDING, new CUDAEmbeddingCreator());
    addCreator(SILU, new CUDASiLUCreator());
    addCreator(F_TTADD, new CUDAAddCreator());
    addCreator(F_TTMUL, new CUDAMulCreator());
    addCreator(F_VIEW, new CUDAViewCreator());
    addCreator(F_CLIP, new CUDAClipCreator());
    addCreator(F_FA2, new CUDAFlashAttention2Creator());
    addCreator(OP_GELU, new CUDAGeluCreator());
    addCreator(QUICKGLUE, new CUDAQuickGeluCreator());
    addCreator(SOFTMAX, new CUDASoftMaxCreator());
    addCreator(F_DIV, new CUDADivScalarCreator());
    addCreator(F_MM, new CUDAMatmulFuncCreator());
    addCreator(F_TRANPOSE, new CUDATransposeCreator());
    addCreator(PARAMETER, new CUDAParameterCreator());
    addCreator(F_CAT, new CUDACatCreator());
    addCreator(F_SPLIT, new CUDASplitCreator());
    addCreator(CONVOLUTION2D, new CUDAConvolution2DCreator());
    addCreator(MULTIMODALROPE, new CUDAMultimodalRoPECreator());
    addCreator(F_APPLY_VISIOROPE, new CUDAVisionRoPEFuncCreator());
}

} // namespace mllm
