// Module.cpp -- LLaMA-family decoder graph on the CUDA backend, op for op the sequence
// models/llama/modeling_llama.hpp:29-115 and models/transformer/modeling_transformer.hpp:132-212
// issue through Layer::run / Tensor::runFunc.
#include "Module.hpp"
#include <cstdlib>
#include <iostream>

namespace mllm {

// CPUBackend::runLayer (CPUBackend.cpp:407-417): "out-<op name>" (or "-i" suffixed for N outputs), then runOp.  The module tag keeps the
// activation names of two loaded models apart.
vector<Tensor> CUDABackend::runLayer(Layer *layer, vector<Tensor> inputs, int N) {
    return runOp(layer->op(), inputs, layer->out_names(N), false);
}


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
    static int instances = 0;
    if (instances++ > 0) st_.tag = "@" + std::to_string(instances);
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
    st_.backend->dropQuantTag();
    CUDABackend *bn = st_.backend;
    last_fused_ = false;
    if (!bn->doTrace && !dump && ids.sequence() == 1 && ids.contiguous() && decode_plan_usable(ids.batch()))
        return decode_step(ids.devPtr<float>(), ids.batch(), false);
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

// ===================================================================== decode plan
struct DecodeLookahead { const void *ptr[8]; size_t bytes[8]; int n = 0; };

struct LlamaModel::DecodePlan {
    struct L {
        FuseInfo n1, n2, q, k, v, o, gate, up, down, rope, kc, vc;
        Op *advance[4]; // q_rope, k_rope, k_cache, v_cache
    };
    int B = 0;
    vector<L> layers;
    FuseInfo embd, norm, head;
    Tensor x, x2, qkv, ao, g, logits, ids, ctl, tok_log, attn_ws;
    // k-quant (Q4_K / Q6_K) Linears have no fused TMA GEMV: they run as rmsnorm -> mllmcu_linear (Q8_K quantiser + k-quant GEMV) ->
    // add / silu*mul launches on these fixed buffers, still inside the one captured graph with the device-resident position
    bool kquant = false;
    Tensor kq_xn, kq_t1, kq_t2, kq_ws;
    size_t kq_ws_bytes = 0;
    // activation replicas: x, x2 and g are stored `reps` times ([reps][B][width]) by their producers and every RMSNorm weight is
    // copied `reps` times ([norm][reps][H]), so the CTAs of the consuming Linear spread their reads over more L2 slices
    int reps = 1;
    Tensor norm_rep;
    size_t attn_ws_bytes = 0;
    int dev_pos = -2;          // value ctl[0] holds in stream order (-2: unknown)
    int ctl_host = 0;          // staging for the H2D of ctl[0]
    bool warmed[2] = {false, false};
    void *graph[2] = {nullptr, nullptr};
    bool unsupported = false, use_graph = true, use_pdl = true;
    int launches = 0;
    void *step = nullptr; // persistent whole-step kernel (mllmcu_decode_step_*), when the shapes fit it
    // L2 look-ahead: the byte ranges of LATER launches' weights that fused Linear number i asks L2 to fetch while it runs
    vector<DecodeLookahead> lookahead;
    bool lookahead_built = false;
    ~DecodePlan() {
        for (void *g : graph) if (g) mllmcu_graph_destroy(g);
        if (step) mllmcu_decode_step_destroy(step);
    }
};

LlamaModel::~LlamaModel() = default;
void LlamaModel::set_step_kernel(bool on) {
    want_step_kernel_ = on;
    st_.backend->sync();
    plan_.reset();
}

static bool env_on(const char *k) {
    const char *e = getenv(k);
    return e && e[0] && e[0] != '0';
}

bool LlamaModel::decode_plan_usable(int B) {
    static const bool eager = env_on("MLLM_CUDA_EAGER_DECODE");
    if (eager || force_eager_ || B < 1 || B > 64) return false; // 1-8: fused TMA GEMVs; 9-64: op-level kernels inside the captured graph
    if (plan_ && plan_->unsupported) return false;
    const int H = cfg_.hidden, D = cfg_.head_dim();
    if (H % 128 || cfg_.ffn % 128 || (D != 32 && D != 64 && D != 128)) return false;
    if (plan_ && plan_->B == B) {
        // the plan (and its captured graph) holds the KV-cache addresses: a prefill at another batch size re-makes the caches
        // (CUDAKVCache::ensure), and coming back to this batch size need not hand back the same blocks
        for (int i = 0; i < cfg_.layers; i++) {
            FuseInfo k, v;
            Block &b = blocks_[i];
            if (!(b.k_cache.op() && b.k_cache.op()->fuseInfo(k) && b.v_cache.op() && b.v_cache.op()->fuseInfo(v)) ||
                k.cache != plan_->layers[i].kc.cache || v.cache != plan_->layers[i].vc.cache) {
                st_.backend->sync();
                plan_.reset();
                break;
            }
        }
    }
    if (!(plan_ && plan_->B == B)) {
        // (re)build: collect the device state of every op on the path
        auto p = std::make_unique<DecodePlan>();
        p->B = B;
        p->layers.resize(cfg_.layers);
        auto get = [](Layer &l, FuseInfo &f) { return l.op() && l.op()->fuseInfo(f); };
        bool ok = get(embedding_, p->embd) && get(norm_, p->norm) && get(lm_head_, p->head);
        for (int i = 0; ok && i < cfg_.layers; i++) {
            Block &b = blocks_[i];
            auto &l = p->layers[i];
            ok = get(b.norm1, l.n1) && get(b.norm2, l.n2) && get(b.q, l.q) && get(b.k, l.k) && get(b.v, l.v) && get(b.o, l.o) &&
                 get(b.gate, l.gate) && get(b.up, l.up) && get(b.down, l.down) && get(b.q_rope, l.rope) && get(b.k_cache, l.kc) &&
                 get(b.v_cache, l.vc);
            if (!ok) break;
            for (const FuseInfo *f : {&l.q, &l.k, &l.v, &l.o, &l.gate, &l.up, &l.down}) {
                ok = ok && (f->wtype == MLLM_TYPE_Q4_0 || f->wtype == MLLM_TYPE_Q4_K || f->wtype == MLLM_TYPE_Q6_K);
                if (f->wtype != MLLM_TYPE_Q4_0) p->kquant = true;
            }
            ok = ok && !l.n1.unit_offset && !l.n2.unit_offset && l.kc.cache_batch == B && l.vc.cache_batch == B && l.kc.cache_dim == D &&
                 l.kc.cache_heads == cfg_.kv_heads && l.rope.rot_dim % 2 == 0;
            l.advance[0] = b.q_rope.op(); l.advance[1] = b.k_rope.op(); l.advance[2] = b.k_cache.op(); l.advance[3] = b.v_cache.op();
        }
        ok = ok && (p->head.wtype == MLLM_TYPE_Q4_0 || p->head.wtype == MLLM_TYPE_Q4_K || p->head.wtype == MLLM_TYPE_Q6_K) &&
             p->embd.wtype == MLLM_TYPE_Q4_0 && !p->norm.unit_offset;
        if (p->head.wtype != MLLM_TYPE_Q4_0) p->kquant = true;
        static const bool no_kq_plan = env_on("MLLM_CUDA_NO_KQUANT_PLAN");
        if (p->kquant && (no_kq_plan || cfg_.hidden % 256 || cfg_.ffn % 256)) ok = false;
        // the fused GEMV holds at most 8 activation rows: 9-32 rows run it in chunks of 8 (later chunks read the layer's weights from L2),
        // larger batches take the op-level kernels (tcgen05 / dp4a Linear)
        if (B > 32 && mllmcu_gemv_fused_max_rows() < B) p->kquant = true;
        if (!ok) return false; // caches not allocated yet (no prefill so far) or a dtype outside the fused kernels
        CUDABackend *bn = st_.backend;
        const int QKV = (cfg_.heads + 2 * cfg_.kv_heads) * D;
        {
            const char *e = getenv("MLLM_CUDA_ACT_REPLICAS");
            p->reps = (e && e[0]) ? atoi(e) : 1; // measured on B200: copies do not shorten the prologue loads (profiles/r01_gemv_timeline.txt)
            if (p->reps < 1) p->reps = 1;
            if (p->reps > 16) p->reps = 16;
            if (B > 8) p->reps = 1; // chunked launches address rows by pointer offsets: one copy of every activation
        }
        const int R = p->reps;
        p->x = Tensor(R, 1, B, H, bn); p->x2 = Tensor(R, 1, B, H, bn);
        p->qkv = Tensor(B, 1, 1, QKV, bn); p->ao = Tensor(B, 1, 1, cfg_.heads * D, bn);
        p->g = Tensor(R, 1, B, cfg_.ffn, bn); p->logits = Tensor(B, 1, 1, cfg_.vocab, bn);
        p->norm_rep = Tensor(2 * cfg_.layers + 1, 1, R, H, bn);
        for (int i = 0; i <= 2 * cfg_.layers; i++) {
            const FuseInfo &n = i == 2 * cfg_.layers ? p->norm : (i % 2 ? p->layers[i / 2].n2 : p->layers[i / 2].n1);
            for (int r = 0; r < R; r++)
                MLLMCU_CHECK(mllmcu_memcpy_d2d(p->norm_rep.devPtr<float>() + ((size_t)i * R + r) * H, n.norm_w, sizeof(float) * H, bn->stream()));
        }
        p->ids = Tensor(B, 1, 1, 1, bn); p->ctl = Tensor(1, 1, 1, 4, bn);
        p->tok_log = Tensor(1, 1, cfg_.cache_limit + 1, B, bn);
        if (p->kquant) {
            p->reps = 1;
            const int W = cfg_.ffn > cfg_.hidden ? cfg_.ffn : cfg_.hidden;
            p->kq_xn = Tensor(1, 1, B, cfg_.hidden, bn);
            p->kq_t1 = Tensor(1, 1, B, W, bn);
            p->kq_t2 = Tensor(1, 1, B, W, bn);
            p->kq_ws_bytes = mllmcu_linear_workspace_bytes(B, W);
            p->kq_ws = Tensor(1, 1, 1, (int)((p->kq_ws_bytes + 3) / 4), bn);
        }
        p->attn_ws_bytes = mllmcu_decode_attn_workspace_bytes(B, cfg_.heads, D);
        p->attn_ws = Tensor(1, 1, 1, (int)((p->attn_ws_bytes + 3) / 4), bn);
        MLLMCU_CHECK(mllmcu_memset(p->attn_ws.rawDevPtr(), 0, p->attn_ws.cntSize(), bn->stream()));
        MLLMCU_CHECK(mllmcu_memset(p->ctl.rawDevPtr(), 0, 16, bn->stream()));
        p->use_graph = !env_on("MLLM_CUDA_NO_GRAPH");
        p->use_pdl = !env_on("MLLM_CUDA_NO_PDL");
        p->launches = 3 + (p->layers[0].rope.rot_dim == D ? 5 : 6) * cfg_.layers;
        // the persistent whole-step kernel is opt-in: on B200 its grid barriers + in-order per-warp rings currently lose
        // to the PDL-chained persistent GEMVs (2.2 vs 1.7 ms/token on the 7B, profiles/r01_step_kernel_phases.txt)
        if ((env_on("MLLM_CUDA_STEP_KERNEL") || want_step_kernel_) && B <= 4 && !p->kquant) build_step_kernel(*p);
        if (p->step) p->launches = 3;
        plan_ = std::move(p);
    }
    // per-step state: every RoPE counter and cache length must agree (they do on the LLaMA path)
    const int len = blocks_[0].k_cache.op()->getCacheSeqLen();
    for (auto &b : blocks_) {
        FuseInfo r;
        if (b.k_cache.op()->getCacheSeqLen() != len || b.v_cache.op()->getCacheSeqLen() != len) return false;
        if (!b.q_rope.op()->fuseInfo(r) || r.pos != len) return false;
        if (!b.k_rope.op()->fuseInfo(r) || r.pos != len) return false;
    }
    return true;
}

int LlamaModel::decode_mode() const {
    if (!plan_ || plan_->unsupported) return 0;
    int m = plan_->step ? 2 : 1;
    if (plan_->graph[0] || plan_->graph[1]) m += 4;
    if (plan_->use_pdl && !plan_->step) m += 8;
    return m;
}
float *LlamaModel::decode_ids() { return plan_ ? plan_->ids.devPtr<float>() : nullptr; }
const int *LlamaModel::decode_tok_log() { return plan_ ? plan_->tok_log.devPtr<int>() : nullptr; }

// describe the step to the persistent kernel: per block qkv (RMSNorm prologue; rope + KV append + attention follow),
// o (+ residual), gate|up (RMSNorm prologue, silu*mul), down (+ residual); then final norm + lm_head
void LlamaModel::build_step_kernel(DecodePlan &p) {
    const int B = p.B, H = cfg_.hidden, D = cfg_.head_dim(), Hq = cfg_.heads, Hkv = cfg_.kv_heads, F = cfg_.ffn;
    const int QKV = (Hq + 2 * Hkv) * D;
    float *x = p.x.devPtr<float>(), *x2 = p.x2.devPtr<float>(), *qkv = p.qkv.devPtr<float>(), *ao = p.ao.devPtr<float>(), *g = p.g.devPtr<float>();
    vector<mllmcu_step_job> jobs;
    vector<void *> kc, vc;
    auto job = [&](std::initializer_list<const FuseInfo *> ws, const float *in, int K, const FuseInfo *norm, int epi, const float *res, int ldr,
                   float *out, int ldy, int post, int layer) {
        mllmcu_step_job j = {};
        int n = 0;
        for (const FuseInfo *f : ws) { j.W[n] = f->weight; j.N[n] = f->out; j.bias[n] = f->bias; n++; }
        j.nW = n; j.K = K; j.epi = epi; j.post = post; j.layer = layer; j.x = in; j.ldx = K;
        j.norm_w = norm ? norm->norm_w : nullptr; j.eps = norm ? norm->eps : 0.f; j.residual = res; j.ldr = ldr; j.y = out; j.ldy = ldy;
        jobs.push_back(j);
    };
    int li = 0;
    for (auto &l : p.layers) {
        job({&l.q, &l.k, &l.v}, x, H, &l.n1, 0, nullptr, 0, qkv, QKV, 1, li);
        job({&l.o}, ao, Hq * D, nullptr, 1, x, H, x2, H, 0, li);
        job({&l.gate, &l.up}, x2, H, &l.n2, 2, nullptr, 0, g, F, 0, li);
        job({&l.down}, g, F, nullptr, 1, x2, H, x, H, 0, li);
        kc.push_back(l.kc.cache);
        vc.push_back(l.vc.cache);
        li++;
    }
    job({&p.head}, x, H, &p.norm, 0, nullptr, 0, p.logits.devPtr<float>(), cfg_.vocab, 0, 0);
    const FuseInfo &r = p.layers[0].rope;
    mllmcu_step_desc d = {};
    d.B = B; d.Hq = Hq; d.Hkv = Hkv; d.D = D; d.rope_type = r.rope_type; d.rot_dim = r.rot_dim; d.tdim = r.rot_dim;
    d.cache_limit = p.layers[0].kc.cache_limit; d.n_layers = cfg_.layers; d.sin_t = r.sin_t; d.cos_t = r.cos_t;
    d.kv_bstride = p.layers[0].kc.cache_bstride; d.kcache = kc.data(); d.vcache = vc.data();
    d.qkv = qkv; d.ldqkv = QKV; d.ao = ao; d.ldao = Hq * D; d.ctl = p.ctl.devPtr<int>();
    st_.backend->sync(); // create() uses synchronous copies on the default stream
    const int rc = mllmcu_decode_step_create(&d, jobs.data(), (int)jobs.size(), &p.step);
    if (rc == -2) { p.step = nullptr; return; } // shapes outside the kernel: keep the chained launches
    MLLMCU_CHECK(rc);
}

// Weights are streamed in a fixed cyclic order (layer 0 q,k,v,o,gate,up,down ... lm_head, then layer 0 again for the next
// token).  Launch i is told to pull into L2 the part of that stream between the previous launch's frontier and
// `lead` bytes past the end of its own weights, so the next launches find their rows in L2 while HBM keeps streaming
// through launch gaps and prologues.  MLLM_CUDA_L2_LEAD_MB sets the lead (default 0 = off: no measurable gain so far).
static void build_lookahead(const vector<vector<std::pair<const void *, size_t>>> &own, size_t lead, vector<DecodeLookahead> &out) {
    struct Seg { const uint8_t *p; size_t bytes, start; };
    vector<Seg> segs;
    vector<size_t> end_of; // stream offset where launch i's own weights end
    size_t total = 0;
    for (auto &launch : own) {
        for (auto &w : launch) { segs.push_back({(const uint8_t *)w.first, w.second, total}); total += w.second; }
        end_of.push_back(total);
    }
    out.assign(own.size(), DecodeLookahead());
    if (lead == 0 || total == 0) return;
    if (total < 4 * lead) lead = (total / 4) & ~(size_t)15; // toy models: never ask for more than a quarter of the stream
    if (lead == 0) return;
    size_t frontier = lead; // absolute position (not wrapped) the previous launch -- the last one of the previous token -- reached
    size_t si = 0;          // segment holding `frontier % total`
    for (size_t i = 0; i < own.size(); i++) {
        const size_t target = end_of[i] + lead;
        DecodeLookahead &la = out[i];
        while (frontier < target && la.n < 8) {
            const size_t pos = frontier % total;
            while (!(segs[si].start <= pos && pos < segs[si].start + segs[si].bytes)) si = (si + 1) % segs.size();
            const size_t off = pos - segs[si].start;
            const size_t take = std::min(segs[si].bytes - off, target - frontier);
            la.ptr[la.n] = segs[si].p + off;
            la.bytes[la.n] = take;
            la.n++;
            frontier += take;
        }
    }
}

// the kernel sequence of one token; returns false if a fused kernel refuses the shape (rc == -2)
bool LlamaModel::issue_step(DecodePlan &p, bool pick, bool pdl_on) {
    CUDABackend *bn = st_.backend;
    void *st = bn->stream();
    const int B = p.B, H = cfg_.hidden, D = cfg_.head_dim(), Hq = cfg_.heads, Hkv = cfg_.kv_heads, F = cfg_.ffn;
    const int QKV = (Hq + 2 * Hkv) * D, pdl = pdl_on ? 1 : 0;
    int *ctl = p.ctl.devPtr<int>();
    float *x = p.x.devPtr<float>(), *x2 = p.x2.devPtr<float>(), *qkv = p.qkv.devPtr<float>(), *ao = p.ao.devPtr<float>(), *g = p.g.devPtr<float>();
    if (!p.lookahead_built) {
        vector<vector<std::pair<const void *, size_t>>> own;
        auto bytes_of = [](const FuseInfo &f) { return (size_t)f.out * (size_t)(f.in / 32) * 18; };
        for (auto &l : p.layers) {
            own.push_back({{l.q.weight, bytes_of(l.q)}, {l.k.weight, bytes_of(l.k)}, {l.v.weight, bytes_of(l.v)}});
            own.push_back({{l.o.weight, bytes_of(l.o)}});
            own.push_back({{l.gate.weight, bytes_of(l.gate)}, {l.up.weight, bytes_of(l.up)}});
            own.push_back({{l.down.weight, bytes_of(l.down)}});
        }
        own.push_back({{p.head.weight, bytes_of(p.head)}});
        const char *e = getenv("MLLM_CUDA_L2_LEAD_MB");
        const size_t lead = (size_t)((e && e[0]) ? atoi(e) : 0) << 20; // off: within run-to-run noise on the 7B (profiles/r01_gemv_timeline.txt)
        build_lookahead(own, lead, p.lookahead);
        p.lookahead_built = true;
    }
    size_t launch_no = 0;
    const int R = p.reps;
    // in_reps / out_reps: how many copies of the input row exist and how many of the output row to store; norm_idx picks the
    // replicated RMSNorm weight (2 * layer, 2 * layer + 1, or 2 * layers for the final norm)
    auto gemv = [&](std::initializer_list<const FuseInfo *> ws, const float *in, int K, int in_reps, const FuseInfo *norm, int norm_idx, int epi,
                    const float *res, int ldr, float *out, int ldy, int out_reps, bool late_trigger = false) -> bool {
        if (p.kquant) { // op-level kernels on the plan's fixed buffers (graph-capturable: no host-side position, no allocation)
            launch_no++;
            const float *src = in;
            if (norm) {
                MLLMCU_CHECK(mllmcu_rmsnorm(in, p.norm_rep.devPtr<float>() + (size_t)norm_idx * R * H, p.kq_xn.devPtr<float>(), B, K, norm->eps, 0, st));
                src = p.kq_xn.devPtr<float>();
            }
            auto lin = [&](const FuseInfo *f, float *y, int ld) {
                if (f->wtype == MLLM_TYPE_Q4_0 && B <= 32) {
                    // 9-32 rows of Q4_0: the decode GEMV in chunks of 8 rows (later chunks read the weights from L2) fills the machine, the
                    // 128-row tcgen05 tile does not (N / 128 CTAs: 16 for a 2048-wide Linear)
                    bool done = true;
                    for (int m0 = 0; m0 < B; m0 += 8) {
                        const void *Wp[1] = {f->weight};
                        const int Ns[1] = {f->out};
                        const float *bs[1] = {f->bias};
                        const int rows = B - m0 < 8 ? B - m0 : 8;
                        const int rc = mllmcu_gemv_fused(MLLM_TYPE_Q4_0, Wp, Ns, bs, 1, src + (size_t)m0 * K, K, rows, K, nullptr, 0.f, 0, nullptr, 0,
                                                         y + (size_t)m0 * ld, ld, 0, 0, st);
                        if (rc == -2) { done = false; break; }
                        MLLMCU_CHECK(rc);
                    }
                    if (done) return;
                }
                MLLMCU_CHECK(mllmcu_linear(f->wtype, f->weight, src, f->bias, y, B, f->out, K, K, ld, 0, p.kq_ws.rawDevPtr(), p.kq_ws_bytes, st));
            };
            if (epi == 0) {
                int col = 0;
                for (const FuseInfo *f : ws) { lin(f, out + col, ldy); col += f->out; }
            } else if (epi == 1) {
                const FuseInfo *f = *ws.begin();
                lin(f, p.kq_t1.devPtr<float>(), f->out);
                if (ldr != f->out || ldy != f->out) throw std::runtime_error("[mllm-cuda] decode plan: residual rows must be dense");
                MLLMCU_CHECK(mllmcu_add(p.kq_t1.devPtr<float>(), res, out, (size_t)B * f->out, st));
            } else {
                const FuseInfo *fg = *ws.begin(), *fu = *(ws.begin() + 1);
                lin(fg, p.kq_t1.devPtr<float>(), fg->out);
                lin(fu, p.kq_t2.devPtr<float>(), fu->out);
                if (ldy != fg->out) throw std::runtime_error("[mllm-cuda] decode plan: gate rows must be dense");
                MLLMCU_CHECK(mllmcu_silu_mul(p.kq_t1.devPtr<float>(), p.kq_t2.devPtr<float>(), out, (size_t)B * fg->out, st));
            }
            return true;
        }
        if (B > 8 && B <= mllmcu_gemv_fused_max_rows()) { // up to 64 rows in ONE launch on the tensor cores (gemv_mma.cu)
            mllmcu_gemv_desc c = {};
            c.wtype = MLLM_TYPE_Q4_0;
            for (const FuseInfo *f : ws) { c.W[c.nW] = f->weight; c.N[c.nW] = f->out; c.bias[c.nW] = f->bias; c.nW++; }
            c.x = in; c.ldx = K; c.M = B; c.K = K;
            if (norm) { c.norm_w = p.norm_rep.devPtr<float>() + (size_t)norm_idx * H; c.eps = norm->eps; c.norm_reps = 1; c.norm_rep_stride = H; }
            c.epilogue = epi; c.residual = res; c.ldr = ldr; c.y = out; c.ldy = ldy;
            c.pdl = (pdl && late_trigger) ? 2 : pdl;
            c.x_reps = 1; c.y_reps = 1;
            const int rc = mllmcu_gemv_fused_ex(&c, st);
            if (rc == 0) { launch_no++; return true; }
            if (rc != -2) MLLMCU_CHECK(rc);
        }
        if (B > 32) return false;
        if (B > 8) { // chunks of 8 rows through the same fused kernel
            launch_no++;
            for (int m0 = 0; m0 < B; m0 += 8) {
                mllmcu_gemv_desc c = {};
                c.wtype = MLLM_TYPE_Q4_0;
                for (const FuseInfo *f : ws) { c.W[c.nW] = f->weight; c.N[c.nW] = f->out; c.bias[c.nW] = f->bias; c.nW++; }
                c.x = in + (size_t)m0 * K; c.ldx = K; c.M = B - m0 < 8 ? B - m0 : 8; c.K = K;
                if (norm) { c.norm_w = p.norm_rep.devPtr<float>() + (size_t)norm_idx * H; c.eps = norm->eps; c.norm_reps = 1; c.norm_rep_stride = H; }
                c.epilogue = epi; c.residual = res ? res + (size_t)m0 * ldr : nullptr; c.ldr = ldr; c.y = out + (size_t)m0 * ldy; c.ldy = ldy;
                c.pdl = (pdl && late_trigger && m0 == 0) ? 2 : pdl;
                c.x_reps = 1; c.y_reps = 1;
                const int rc = mllmcu_gemv_fused_ex(&c, st);
                if (rc == -2) return false;
                MLLMCU_CHECK(rc);
            }
            return true;
        }
        mllmcu_gemv_desc d = {};
        d.wtype = MLLM_TYPE_Q4_0;
        for (const FuseInfo *f : ws) { d.W[d.nW] = f->weight; d.N[d.nW] = f->out; d.bias[d.nW] = f->bias; d.nW++; }
        d.x = in; d.ldx = K; d.M = B; d.K = K;
        if (norm) {
            d.norm_w = p.norm_rep.devPtr<float>() + (size_t)norm_idx * R * H;
            d.eps = norm->eps; d.norm_reps = R; d.norm_rep_stride = H;
        }
        d.epilogue = epi; d.residual = res; d.ldr = ldr; d.y = out; d.ldy = ldy; d.pdl = (pdl && late_trigger) ? 2 : pdl;
        d.x_reps = in_reps; d.x_rep_stride = (long long)B * K;
        d.y_reps = out_reps; d.y_rep_stride = (long long)B * ldy;
        const DecodeLookahead &la = p.lookahead[launch_no++];
        d.n_pf = la.n;
        for (int i = 0; i < la.n; i++) { d.pf_ptr[i] = la.ptr[i]; d.pf_bytes[i] = la.bytes[i]; }
        const int rc = mllmcu_gemv_fused_ex(&d, st);
        if (rc == -2) return false;
        MLLMCU_CHECK(rc);
        return true;
    };
    MLLMCU_CHECK(mllmcu_decode_embed(p.embd.weight, p.ids.devPtr<float>(), x, B, H, ctl, st));
    if (p.step) {
        MLLMCU_CHECK(mllmcu_decode_step_launch(p.step, st));
        if (pick)
            MLLMCU_CHECK(mllmcu_decode_argmax(p.logits.devPtr<float>(), B, cfg_.vocab, p.ids.devPtr<float>(), p.tok_log.devPtr<int>(), ctl,
                                              cfg_.cache_limit + 1, st));
        return true;
    }
    for (auto &l : p.layers) {
        const int li = (int)(&l - &p.layers[0]);
        // Layer 0's projection releases its dependents only after its own wait: from then on every kernel of the step starts
        // with the embedding kernel (which bumps ctl[0]) complete, so the attention kernels may read the position, their K/V
        // rows and the RoPE table row ahead of THEIR waits (pdl == 2 below).
        if (!gemv({&l.q, &l.k, &l.v}, x, H, li == 0 ? 1 : R, &l.n1, 2 * li, 0, nullptr, 0, qkv, QKV, 1, li == 0)) return false; // the embedding wrote one copy
        static const bool no_rope_fusion = env_on("MLLM_CUDA_NO_ROPE_FUSION");
        if (l.rope.rot_dim == D && !no_rope_fusion) { // full rotary: RoPE and the KV append ride inside the attention kernel
            MLLMCU_CHECK(mllmcu_decode_attn_rope(l.rope.rope_type, qkv, QKV, l.kc.cache, l.vc.cache, ao, Hq * D, ctl, B, Hq, Hkv, D, l.rope.sin_t,
                                                 l.rope.cos_t, l.rope.rot_dim, l.kc.cache_bstride, l.kc.cache_limit, p.attn_ws.rawDevPtr(),
                                                 p.attn_ws_bytes, (pdl && !p.kquant) ? 2 : 0, st));
        } else {
            MLLMCU_CHECK(mllmcu_decode_rope_kv(l.rope.rope_type, qkv, QKV, B, Hq, Hkv, D, l.rope.rot_dim, l.rope.sin_t, l.rope.cos_t, l.rope.rot_dim,
                                               ctl, l.kc.cache, l.vc.cache, l.kc.cache_bstride, l.kc.cache_limit, pdl, st));
            MLLMCU_CHECK(mllmcu_decode_attn(qkv, QKV, l.kc.cache, l.vc.cache, ao, Hq * D, ctl, B, Hq, Hkv, D, l.kc.cache_bstride,
                                            p.attn_ws.rawDevPtr(), p.attn_ws_bytes, pdl, st));
        }
        if (!gemv({&l.o}, ao, Hq * D, 1, nullptr, 0, 1, x, H, x2, H, R)) return false;
        if (!gemv({&l.gate, &l.up}, x2, H, R, &l.n2, 2 * li + 1, 2, nullptr, 0, g, F, R)) return false;
        if (!gemv({&l.down}, g, F, R, nullptr, 0, 1, x2, H, x, H, R)) return false;
    }
    if (!gemv({&p.head}, x, H, R, &p.norm, 2 * cfg_.layers, 0, nullptr, 0, p.logits.devPtr<float>(), cfg_.vocab, 1)) return false;
    if (pick)
        MLLMCU_CHECK(mllmcu_decode_argmax(p.logits.devPtr<float>(), B, cfg_.vocab, p.ids.devPtr<float>(), p.tok_log.devPtr<int>(), ctl,
                                          cfg_.cache_limit + 1, st));
    return true;
}

int LlamaModel::cache_seq_len() { return blocks_[0].k_cache.op() ? blocks_[0].k_cache.op()->getCacheSeqLen() : 0; }

Tensor LlamaModel::decode_step(const float *ids_dev, int B, bool pick) {
    st_.backend->dropQuantTag(); // the plan's kernels write outside runOp
    CUDABackend *bn = st_.backend;
    DecodePlan &p = *plan_;
    void *st = bn->stream();
    const int len = blocks_[0].k_cache.op()->getCacheSeqLen();
    if (len + 1 > cfg_.cache_limit) { // same refusal as CUDAKVCache::reshape (CPUKVCache.cpp:121-126)
        std::cerr << "\n[ERROR]: Current tokens exceed cache limit: " << len + 1 << ">" << cfg_.cache_limit << ";"
                  << "\n         Please set args `--limits` >" << cfg_.cache_limit << std::endl;
        throw std::runtime_error("[mllm-cuda] KV cache limit exceeded");
    }
    if (len + 1 > p.layers[0].rope.pos_max) throw std::runtime_error("[mllm-cuda] RoPE: position exceeds max_position_embeddings");
    if (ids_dev != p.ids.devPtr<float>()) MLLMCU_CHECK(mllmcu_memcpy_d2d(p.ids.rawDevPtr(), ids_dev, sizeof(float) * B, st));
    if (p.dev_pos != len - 1) { // the embed kernel pre-increments
        p.ctl_host = len - 1;
        MLLMCU_CHECK(mllmcu_memcpy_h2d(p.ctl.rawDevPtr(), &p.ctl_host, sizeof(int), st));
        p.dev_pos = len - 1;
    }
    const int gi = pick ? 1 : 0;
    bool ran = false;
    if (p.use_graph && !p.graph[gi] && p.warmed[gi]) {
        // second step of this flavour: every kernel attribute is configured, capture the sequence
        for (int attempt = 0; attempt < 2 && !p.graph[gi]; attempt++) {
            MLLMCU_CHECK(mllmcu_graph_begin_capture(st));
            bool ok = false;
            try { ok = issue_step(p, pick, p.use_pdl); } catch (...) { mllmcu_graph_abort_capture(st); throw; }
            void *ge = nullptr;
            if (!ok) { mllmcu_graph_abort_capture(st); break; }
            if (mllmcu_graph_end_capture(st, &ge) == 0) p.graph[gi] = ge;
            else if (p.use_pdl) p.use_pdl = false; // retry the capture with plain stream order
            else { p.use_graph = false; break; }
        }
    }
    if (p.graph[gi]) {
        MLLMCU_CHECK(mllmcu_graph_launch(p.graph[gi], st));
        ran = true;
    } else {
        ran = issue_step(p, pick, p.use_pdl);
        p.warmed[gi] = true;
    }
    if (!ran) { // a kernel refused the shape mid-step: forget the plan and redo the token op by op
        p.unsupported = true;
        p.dev_pos = -2;
        Tensor ids(B, 1, 1, 1, bn);
        ids.setName("input_ids");
        MLLMCU_CHECK(mllmcu_memcpy_d2d(ids.rawDevPtr(), p.ids.rawDevPtr(), sizeof(float) * B, st));
        Tensor lg = forward(ids);
        if (pick) { // the caller still expects the pick in decode_ids() and in the token log under the consumed position
            p.ctl_host = len;
            MLLMCU_CHECK(mllmcu_memcpy_h2d(p.ctl.rawDevPtr(), &p.ctl_host, sizeof(int), st));
            MLLMCU_CHECK(mllmcu_decode_argmax(lg.devPtr<float>(), B, cfg_.vocab, p.ids.devPtr<float>(), p.tok_log.devPtr<int>(),
                                              p.ctl.devPtr<int>(), cfg_.cache_limit + 1, st));
        }
        return lg;
    }
    p.dev_pos = len;
    for (auto &l : p.layers)
        for (Op *o : l.advance) o->advance(1);
    bn->kernel_launches += p.launches - (pick ? 0 : 1);
    last_fused_ = true;
    return p.logits;
}

} // namespace mllm
