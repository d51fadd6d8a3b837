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
        if (dt != MLLM_TYPE_Q4_0 && dt != MLLM_TYPE_Q8_0)
            throw std::runtime_error("[mllm-cuda] Linear " + name() + ": weight dtype " + std::to_string((int)dt) + " not supported by the CUDA backend yet");
        weight_.setDtype(dt);
        weight_.alloc();
        if (!absent) loader.load(&weight_);
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
        // one call when both sides are dense over (batch, seq); else one call per batch row-block
        const bool dense = x.contiguous() && (y.batch() == 1 || y.batchStride() == (long long)y.sequence() * y.seqStride());
        const int nb = dense ? 1 : x.batch();
        const int M = dense ? x.batch() * x.sequence() : x.sequence();
        const size_t wsb = mllmcu_linear_workspace_bytes(M, in_);
        void *ws = b->workspace(wsb);
        for (int i = 0; i < nb; i++) {
            const float *xp = x.devPtr<float>() + (size_t)i * x.batchStride();
            char *yp = (char *)y.rawDevPtr() + (size_t)i * y.batchStride() * esz;
            MLLMCU_CHECK(mllmcu_linear((int)weight_.dtype(), weight_.rawDevPtr(), xp, bias, yp, M, out_, in_, (int)x.seqStride(),
                                       (int)y.seqStride(), f16, ws, b->workspaceBytes(), b->stream()));
            b->kernel_launches += 2;
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
        if ((int)getp(p, "rope_type", 0) != 0) throw std::runtime_error("[mllm-cuda] RoPE " + name + ": rope scaling not supported yet");
    }
    ErrorCode reshape(vector<shared_ptr<Tensor>> in, vector<shared_ptr<Tensor>> out) override {
        out[0]->reshape(in[0]->batch(), in[0]->head(), in[0]->sequence(), in[0]->dimension());
        ishape_ = (int)(in[0]->dimension() * partial_);
        char key[128];
        snprintf(key, sizeof key, "%d-%d-%.9g-%d", pose_type_, ishape_, rope_theta_, pos_max_);
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
        if (dt != MLLM_TYPE_Q4_0) throw std::runtime_error("[mllm-cuda] Embedding " + name() + ": only Q4_0 tables (the quantizer forces embed_tokens to Q4_0)");
        weight_.setDtype(dt);
        weight_.alloc();
        if (!absent) loader.load(&weight_);
        return MLLM_NO_ERROR;
    }
    ErrorCode execute(vector<shared_ptr<Tensor>> in, vector<shared_ptr<Tensor>> out) override {
        auto *b = cb(backend_);
        const int n = in[0]->batch() * in[0]->sequence();
        MLLMCU_CHECK(mllmcu_memset(out[0]->rawDevPtr(), 0, out[0]->cntSize(), b->stream()));
        MLLMCU_CHECK(mllmcu_embedding_q4_0(weight_.rawDevPtr(), in[0]->devPtr<float>(), out[0]->devPtr<float>(), n, hidden_, b->stream()));
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
        if (k.dtype() != MLLM_TYPE_F16 || v.dtype() != MLLM_TYPE_F16) throw std::runtime_error("[mllm-cuda] flash_attention2: fp16 K/V (KV cache) expected");
        if (k.seqStride() != (long long)k.head() * k.dimension()) throw std::runtime_error("[mllm-cuda] flash_attention2: K/V rows must be dense");
        const size_t wsb = mllmcu_flash_attn_workspace_bytes(q.batch(), q.sequence(), q.head(), q.dimension());
        void *ws = b->workspace(wsb);
        MLLMCU_CHECK(mllmcu_flash_attn_f16kv(q.devPtr<float>(), k.rawDevPtr(), v.rawDevPtr(), out[0]->devPtr<float>(), q.batch(), q.sequence(), k.sequence(),
                                             q.head(), k.head(), q.dimension(), causal_, k.batchStride(), ws, b->workspaceBytes(), b->stream()));
        b->kernel_launches += 2;
        return MLLM_NO_ERROR;
    }

private:
    bool causal_;
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
}

} // namespace mllm
