// Module.cpp -- LLaMA-family decoder graph on the CUDA backend, op for op the sequence
// models/llama/modeling_llama.hpp:29-115 and models/transformer/modeling_transformer.hpp:132-212
// issue through Layer::run / Tensor::runFunc.
#include "Module.hpp"

namespace mllm {

LlamaNames LlamaNames::make(int rope_type) {
    LlamaNames n;
    if (rope_type == LLAMAROPE) {
        n = {"layers.", "attention.", "feed_forward.", "wq", "wk", "wv", "wo", "w1", "w3", "w2", "attention_norm", "ffn_norm",
             "tok_embeddings", "norm", "output"};
    } else if (rope_type == HFHUBROPE) {
        n = {"model.layers.", "self_attn.", "mlp.", "q_proj", "k_proj", "v_proj", "o_proj", "gate_proj", "up_proj", "down_proj",
             "input_layernorm", "post_attention_layernorm", "model.embed_tokens", "model.norm", "lm_head"};
    } else {
        throw std::runtime_error("Unsupported llama type");
    }
    return n;
}

static OpParam P(std::initializer_list<std::pair<const string, float>> l) { return OpParam(l); }

LlamaModel::LlamaModel(CUDABackend *bn, const LlamaConfig &cfg) : cfg_(cfg), fn_(&st_) {
    st_.backend = bn;
    const LlamaNames nm = LlamaNames::make(cfg.rope_type);
    const int H = cfg.hidden, D = cfg.head_dim();
    embedding_ = Layer(&st_, EMBEDDING, nm.embd, P({{"hidden_size", (float)H}, {"vocab_size", (float)cfg.vocab}}));
    norm_ = Layer(&st_, RMSNORM, nm.post_norm, P({{"norm_size", (float)H}, {"epsilon", cfg.rms_eps}}));
    lm_head_ = Layer(&st_, LINEAR, nm.lm_head, P({{"in_features", (float)H}, {"out_features", (float)cfg.vocab}, {"bias", 0}}));
    blocks_.resize(cfg.layers);
    for (int i = 0; i < cfg.layers; i++) {
        Block &b = blocks_[i];
        const string base = nm.blk + std::to_string(i) + ".";
        const string a = base + nm.attn, f = base + nm.ffn;
        auto lin = [&](const string &name, int in, int out, bool bias) {
            return Layer(&st_, LINEAR, name, P({{"in_features", (float)in}, {"out_features", (float)out}, {"bias", bias ? 1.f : 0.f}}));
        };
        b.norm1 = Layer(&st_, RMSNORM, base + nm.attn_norm, P({{"norm_size", (float)H}, {"epsilon", cfg.rms_eps}}));
        b.norm2 = Layer(&st_, RMSNORM, base + nm.ffn_norm, P({{"norm_size", (float)H}, {"epsilon", cfg.rms_eps}}));
        b.q = lin(a + nm.q, H, cfg.heads * D, cfg.qkv_bias);
        b.k = lin(a + nm.k, H, cfg.kv_heads * D, cfg.qkv_bias);
        b.v = lin(a + nm.v, H, cfg.kv_heads * D, cfg.qkv_bias);
        b.o = lin(a + nm.o, cfg.heads * D, H, false);
        OpParam rp = P({{"pose_type", (float)cfg.rope_type}, {"rope_theta", cfg.rope_theta}, {"max_position_embeddings", (float)cfg.max_pos},
                        {"partial_rotary_factor", 1.0f}, {"rope_type", 0}});
        b.q_rope = Layer(&st_, ROPE, a + "q_rope", rp);
        b.k_rope = Layer(&st_, ROPE, a + "k_rope", rp);
        OpParam kp = P({{"head", (float)cfg.kv_heads}, {"hidden", (float)D}, {"n_rep", (float)(cfg.heads / cfg.kv_heads)},
                        {"cache_max", (float)cfg.cache_limit}, {"for_xnn", 0}, {"fa2", 1}});
        b.k_cache = Layer(&st_, KVCACHE, a + "k_cache", kp);
        b.v_cache = Layer(&st_, KVCACHE, a + "v_cache", kp);
        b.gate = lin(f + nm.gate, H, cfg.ffn, false);
        b.up = lin(f + nm.up, H, cfg.ffn, false);
        b.down = lin(f + nm.down, cfg.ffn, H, false);
        b.silu = Layer(&st_, SILU, f + "act");
    }
}

Tensor LlamaModel::block_forward(Block &b, int idx, Tensor in) {
    const int Hq = cfg_.heads, Hkv = cfg_.kv_heads, D = cfg_.head_dim();
    auto view = [&](Tensor t, int h, int d) {
        return fn_.run(F_VIEW, t.name() + "-view", P({{"b", -1}, {"h", (float)h}, {"s", -1}, {"d", (float)d}}), {t});
    };
    // LLaMABlock::Forward (modeling_llama.hpp:60-68)
    Tensor x = b.norm1(in);
    // MultiHeadAttention::Forward (modeling_transformer.hpp:132-212)
    Tensor q = b.q(x), k = b.k(x), v = b.v(x);
    q = view(q, Hq, D);
    k = view(k, Hkv, D);
    v = view(v, Hkv, D);
    q = b.q_rope(q);
    k = b.k_rope(k);
    k = b.k_cache(k);
    v = b.v_cache(v);
    Tensor o = fn_.run(F_FA2, q.name() + "-" + k.name() + "-fa2", P({{"causal_mask", 1.0f}}), {q, k, v});
    o = view(o, 1, D * Hq);
    o = b.o(o);
    Tensor tmp = fn_.run(F_TTADD, o.name() + "-TTadd", {}, {o, in});
    x = b.norm2(tmp);
    // LLaMAMLP::Forward (modeling_llama.hpp:29-36)
    Tensor g = b.gate(x);
    g = b.silu(g);
    Tensor y = b.up(x);
    g = fn_.run(F_TTMUL, g.name() + "-TTmul", {}, {g, y});
    g = b.down(g);
    return fn_.run(F_TTADD, g.name() + "-TTadd", {}, {g, tmp});
}

Tensor LlamaModel::forward(Tensor ids, float *dump) {
    CUDABackend *bn = st_.backend;
    Tensor x = embedding_(ids);
    const size_t hs = x.cntSize();
    if (dump && !bn->doTrace) MLLMCU_CHECK(mllmcu_memcpy_d2d(dump, x.rawDevPtr(), hs, bn->stream()));
    for (int i = 0; i < cfg_.layers; i++) {
        x = block_forward(blocks_[i], i, x);
        if (dump && !bn->doTrace) MLLMCU_CHECK(mllmcu_memcpy_d2d((char *)dump + (size_t)(i + 1) * hs, x.rawDevPtr(), hs, bn->stream()));
    }
    x = norm_(x);
    if (x.sequence() > 1) x = fn_.run(F_CLIP, x.name() + "-clip-", P({{"b_size", 0}, {"h_size", 0}, {"s_size", 1}, {"d_size", 0}, {"s_0", -1}}), {x});
    return lm_head_(x);
}

void LlamaModel::load(const string &path) {
    loader_ = std::make_unique<ParamLoader>(path);
    st_.loader = loader_.get();
    st_.doLoad = true;
    st_.backend->doTrace = true; // Module::load -> forwardNoInput: create ops, load weights, setUp only (Module.hpp:103-140,215-225)
    Tensor dummy(1, 1, 2, 1, st_.backend);
    forward(dummy);
    st_.backend->doTrace = false;
    st_.doLoad = false;
    st_.backend->sync();
}

void LlamaModel::clear_kvcache() { // LLaMAModel::clear_kvcache (modeling_llama.hpp:117-124)
    for (auto &b : blocks_) {
        for (Layer *l : {&b.k_cache, &b.v_cache, &b.q_rope, &b.k_rope})
            if (l->op()) l->op()->clearCache();
    }
}

} // namespace mllm
