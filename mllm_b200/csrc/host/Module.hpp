// Module.hpp -- the thin slice of mllm's eager front-end the hot path is entered through
// (mllm/Layer.hpp:116-222 Layer::run / initOp, mllm/Module.hpp:193-261 load / operator()),
// and the LLaMA-family graph exactly as the reference's model headers compose it
// (models/llama/modeling_llama.hpp:29-115, models/transformer/modeling_transformer.hpp:132-212).
// Inside the mllm tree none of this is needed: the unchanged model headers drive the
// CUDABackend through Backend::runOp.  Here it lets the same op sequence run standalone.
#pragma once
#include "mllm_cuda.hpp"

namespace mllm {

struct ModuleState { // Module::doLoad / doTrace / loader (mllm/Module.hpp)
    CUDABackend *backend = nullptr;
    ParamLoader *loader = nullptr;
    bool doLoad = false;
    string tag; // suffix making this module's activation names unique inside the backend (two models may be loaded at once)
};

class Layer { // mllm/Layer.hpp:30-222
public:
    Layer() = default;
    Layer(ModuleState *st, OpType type, const string &name, OpParam p = {}) : st_(st), name_(name), param_(std::move(p)) { param_["type"] = (float)type; }
    bool ready() const { return st_ != nullptr; }
    Tensor operator()(Tensor x) { return run({x})[0]; }
    vector<Tensor> run(vector<Tensor> inputs) {
        if (!op_) op_.reset(st_->backend->opCreate(param_, name_)); // part 2: initOp
        if (st_->doLoad && !loaded_) {                                // part 3: load params
            op_->load(*st_->loader);
            loaded_ = true;
        }
        return st_->backend->runLayer(this, inputs, 1); // part 4 (Backend::runLayer names the outputs and calls runOp)
    }
    Op *op() { return op_.get(); }
    const string &name() const { return name_; }
    const string &tag() const { return st_->tag; }
    // "out-<op>[-i]<tag>", built once per layer (runLayer is on the host's critical path for short prompts on small models)
    const vector<string> &out_names(int N) {
        const int count = N > 1 ? N : 1;
        if ((int)out_names_.size() != count) {
            out_names_.clear();
            for (int i = 0; i < count; i++) out_names_.push_back("out-" + name_ + (N > 1 ? "-" + std::to_string(i) : "") + st_->tag);
        }
        return out_names_;
    }

private:
    ModuleState *st_ = nullptr;
    string name_;
    OpParam param_;
    std::unique_ptr<Op> op_;
    bool loaded_ = false;
    vector<string> out_names_;
};

// Tensor::runFunc-style functional ops (mllm/Tensor.cpp:361-422): ops created per call site, cached by name.
class FuncCache {
public:
    explicit FuncCache(ModuleState *st) : st_(st) {}
    Tensor run(OpType t, const string &out_name, OpParam p, vector<Tensor> in) {
        auto it = ops_.find(out_name);
        if (it == ops_.end()) {
            p["type"] = (float)t;
            it = ops_.emplace(out_name, std::unique_ptr<Op>(st_->backend->opCreate(p, out_name))).first;
        }
        return st_->backend->runOp(it->second.get(), in, {out_name}, false)[0];
    }

private:
    ModuleState *st_;
    map<string, std::unique_ptr<Op>> ops_;
};

struct LlamaNames { // LLaMANameConfig::init (models/llama/configuration_llama.hpp:20-60)
    string blk, attn, ffn, q, k, v, o, gate, up, down, attn_norm, ffn_norm, embd, post_norm, lm_head;
    static LlamaNames make(int rope_type);
};

struct LlamaConfig {
    int vocab = 32000, hidden = 4096, heads = 32, kv_heads = 32, ffn = 11008, layers = 32;
    int rope_type = LLAMAROPE, cache_limit = 1024, max_pos = 16384;
    float rope_theta = 10000.f, rms_eps = 1e-6f;
    bool qkv_bias = false;
    int head_dim() const { return hidden / heads; }
};

class LlamaModel {
public:
    LlamaModel(CUDABackend *bn, const LlamaConfig &cfg);
    void load(const string &path); // Module::load: trace pass with doLoad, then ready
    // ids: device fp32 [B,1,S,1]; returns logits [B,1,1,vocab] (device).  dump (optional, device
    // fp32 [(layers+1)*B*S*hidden]) receives the hidden state after the embedding and each block.
    Tensor forward(Tensor ids, float *dump = nullptr);
    void clear_kvcache();
    // one transformer block alone on a given hidden state [B,1,S,hidden] (parity tests feed it the reference's own layer input)
    Tensor run_block(int layer, Tensor x) { return block_forward(blocks_.at(layer), layer, x); }
    const LlamaConfig &config() const { return cfg_; }
    CUDABackend *backend() { return st_.backend; }
    ~LlamaModel();

    // ---- fused single-token step (Module.cpp, "decode plan"): the S == 1 forward as ~6 launches per block
    // (qkv GEMV with RMSNorm prologue, RoPE + KV append, attention, o GEMV + residual, gate/up GEMV with
    // RMSNorm prologue and silu*mul epilogue, down GEMV + residual), captured once in a CUDA graph.
    // forward() takes it automatically when the shapes qualify; MLLM_CUDA_EAGER_DECODE=1 forces op-by-op.
    bool decode_plan_usable(int B);
    Tensor decode_step(const float *ids_dev, int B, bool pick_on_device);
    float *decode_ids();         // device fp32 [B]: input of the next step (written by the on-device pick)
    const int *decode_tok_log(); // device int32 [cache_limit][B]: token picked after consuming position p
    bool last_step_fused() const { return last_fused_; }
    int cache_seq_len();         // tokens already in the KV cache (CPUKVCache::getCacheSeqLen of the first block)
    void set_eager_decode(bool on) { force_eager_ = on; }
    void set_step_kernel(bool on); // persistent whole-step kernel (opt-in); drops the current plan
    // 0: no plan, 1: chained fused kernels, 2: persistent step kernel; +4 when replayed from a CUDA graph, +8 with PDL edges
    int decode_mode() const;

private:
    struct DecodePlan;
    std::unique_ptr<DecodePlan> plan_;
    bool last_fused_ = false, force_eager_ = false, want_step_kernel_ = false;
    bool issue_step(DecodePlan &p, bool pick, bool pdl);
    void build_step_kernel(DecodePlan &p);
    struct Block {
        Layer norm1, norm2, q, k, v, o, q_rope, k_rope, k_cache, v_cache, gate, up, down, silu;
    };
    Tensor block_forward(Block &b, int idx, Tensor x);
    LlamaConfig cfg_;
    ModuleState st_;
    FuncCache fn_;
    Layer embedding_, norm_, lm_head_;
    vector<Block> blocks_;
    std::unique_ptr<ParamLoader> loader_;
};

} // namespace mllm
