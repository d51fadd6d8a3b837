// capi_host.cpp -- C entry points over the host-side backend, so tests and bench.py can
// drive CUDABackend / ops / LlamaModel through ctypes with HOST buffers, the way an mllm
// demo drives a Module (examples/demo_llama.cpp, examples/mllm_benchmark.cpp:53-78,
// Module::generate mllm/Module.cpp:63-205).  Errors: C++ exceptions are caught, the message
// is kept for mllm_cuda_last_error(), and a negative code is returned.
#include "Module.hpp"
#include <cstdlib>

#include <chrono>
#include <cstring>

using namespace mllm;

#define HAPI extern "C" __attribute__((visibility("default")))

static thread_local std::string g_err;
#define GUARD(...)                           \
    try {                                    \
        __VA_ARGS__;                         \
        return 0;                            \
    } catch (const std::exception &e) {      \
        g_err = e.what();                    \
        return -1;                           \
    }

struct HostModel {
    std::unique_ptr<LlamaModel> model;
    double last_ms = 0;
    void *pin = nullptr; // pinned staging for the request's ids (H2D) and token ids (D2H)
    size_t pin_bytes = 0;
    void *pinned(size_t bytes) {
        if (bytes > pin_bytes) {
            if (pin) mllmcu_host_free(pin);
            pin = nullptr;
            pin_bytes = (bytes + 4095) / 4096 * 4096;
            MLLMCU_CHECK(mllmcu_host_alloc(&pin, pin_bytes));
        }
        return pin;
    }
    ~HostModel() { if (pin) mllmcu_host_free(pin); }
};

HAPI const char *mllm_cuda_last_error() { return g_err.c_str(); }

// Backend::global_backends[MLLM_CUDA] = make_unique<CUDABackend>() -- the `case MLLM_CUDA:` of
// Module::initBackend (mllm/Module.hpp:151-190)
HAPI int mllm_cuda_init_backend(int device) {
    GUARD({
        if (!Backend::global_backends.count(MLLM_CUDA)) Backend::global_backends[MLLM_CUDA] = std::make_unique<CUDABackend>(device);
    })
}
static CUDABackend *backend() {
    auto it = Backend::global_backends.find(MLLM_CUDA);
    if (it == Backend::global_backends.end()) throw std::runtime_error("[mllm-cuda] backend not initialised (mllm_cuda_init_backend)");
    return static_cast<CUDABackend *>(it->second.get());
}
HAPI int mllm_cuda_shutdown() { GUARD({ Backend::global_backends.erase(MLLM_CUDA); }) }

// cfg_i: vocab, hidden, layers, heads, kv_heads, ffn, rope_type, cache_limit, max_pos, qkv_bias ; cfg_f: rope_theta, rms_eps
HAPI int mllm_cuda_llama_create(const int *ci, const float *cf, void **out) {
    GUARD({
        LlamaConfig c;
        c.vocab = ci[0]; c.hidden = ci[1]; c.layers = ci[2]; c.heads = ci[3]; c.kv_heads = ci[4]; c.ffn = ci[5];
        c.rope_type = ci[6]; c.cache_limit = ci[7]; c.max_pos = ci[8]; c.qkv_bias = ci[9] != 0;
        c.rope_theta = cf[0]; c.rms_eps = cf[1];
        auto *h = new HostModel();
        h->model = std::make_unique<LlamaModel>(backend(), c);
        *out = h;
    })
}
HAPI int mllm_cuda_llama_destroy(void *h) { GUARD({ delete (HostModel *)h; }) }
HAPI int mllm_cuda_llama_load(void *h, const char *path) { GUARD({ ((HostModel *)h)->model->load(path); }) }
HAPI int mllm_cuda_llama_clear_kvcache(void *h) { GUARD({ ((HostModel *)h)->model->clear_kvcache(); }) }

// model({input_ids}) with host buffers: ids_host [B,S] fp32 token ids -> logits_host [B,vocab].
// dump_host (optional) [(layers+1), B*S, hidden].  Wall time of H2D + forward + D2H is recorded.
HAPI int mllm_cuda_llama_forward(void *hv, const float *ids_host, int B, int S, float *logits_host, float *dump_host) {
    GUARD({
        auto *h = (HostModel *)hv;
        CUDABackend *bn = h->model->backend();
        const auto t0 = std::chrono::steady_clock::now();
        Tensor ids(B, 1, S, 1, bn);
        ids.setName("input_ids");
        ids.copyFromHost(ids_host);
        Tensor dump;
        const LlamaConfig &c = h->model->config();
        if (dump_host) dump = Tensor(1, 1, (c.layers + 1) * B * S, c.hidden, bn);
        // Module::operator() -> Backend::runForward (mllm/Module.hpp:193-261): the outermost forward is timed, device drained before the clock stops
        Tensor logits = bn->runForward([&]() { return vector<Tensor>{h->model->forward(ids, dump_host ? dump.devPtr<float>() : nullptr)}; }, &h->last_ms)[0];
        logits.copyToHost(logits_host);
        if (dump_host) dump.copyToHost(dump_host);
        (void)t0;
    })
}

// Teacher-forced block: x_host [B][S][hidden] is uploaded, block `layer` (norm, q/k/v, RoPE, KV append, attention, o, residual, norm, MLP,
// residual: LLaMABlock::Forward, models/llama/modeling_llama.hpp:60-68) runs alone on an EMPTY KV cache, the result comes back.
// Parity tests feed it the reference's own input of that layer, so each layer is compared without upstream differences.
HAPI int mllm_cuda_llama_block_forward(void *hv, int layer, const float *x_host, int B, int S, float *y_host) {
    GUARD({
        auto *h = (HostModel *)hv;
        CUDABackend *bn = h->model->backend();
        h->model->clear_kvcache();
        Tensor x(B, 1, S, h->model->config().hidden, bn);
        x.setName("teacher_forced_x");
        x.copyFromHost(x_host);
        Tensor y = h->model->run_block(layer, x);
        y.copyToHost(y_host);
        h->model->clear_kvcache();
    })
}

// Greedy generation, the loop of Module::generate (mllm/Module.cpp:63-100,146-205): forward,
// D2H logits, host argmax (LlmTextGenerator greedy, Generate.cpp:46-53), feed the token back.
// ids_host [B,S]; out_tokens [B,max_new] int32.
HAPI int mllm_cuda_llama_request_device(void *hv, const float *ids_host, int B, int S, int n_decode, float *ms_prefill,
                                        float *ms_decode, int *tokens_out);
// Module::generate with the greedy LlmTextGenerator (mllm/Module.cpp:63-205, Generate.cpp:46-53): prompt ids go to the device
// once, every pick is made on the device (first maximum wins, as std::max_element) and feeds the next step without leaving HBM,
// the max_new token ids come back in one copy.  MLLM_CUDA_HOST_PICK=1 keeps the reference's literal loop (logits to the host
// every step, host argmax) for comparison.
HAPI int mllm_cuda_llama_generate(void *hv, const float *ids_host, int B, int S, int max_new, int *out_tokens) {
    static const bool host_pick = [] { const char *e = getenv("MLLM_CUDA_HOST_PICK"); return e && e[0] == '1'; }();
    if (!host_pick) {
        float a = 0.f, b = 0.f;
        return mllm_cuda_llama_request_device(hv, ids_host, B, S, max_new, &a, &b, out_tokens);
    }
    GUARD({
        auto *h = (HostModel *)hv;
        CUDABackend *bn = h->model->backend();
        const int V = h->model->config().vocab;
        std::vector<float> logits((size_t)B * V), next((size_t)B);
        std::vector<float> cur(ids_host, ids_host + (size_t)B * S);
        int s = S;
        for (int step = 0; step < max_new; step++) {
            Tensor ids(B, 1, s, 1, bn);
            ids.setName("input_ids");
            ids.copyFromHost(cur.data());
            Tensor lg = h->model->forward(ids);
            lg.copyToHost(logits.data());
            for (int b = 0; b < B; b++) {
                const float *row = logits.data() + (size_t)b * V;
                int best = 0;
                for (int i = 1; i < V; i++) if (row[i] > row[best]) best = i;
                out_tokens[(size_t)b * max_new + step] = best;
                next[b] = (float)best;
            }
            cur.assign(next.begin(), next.end());
            s = 1;
        }
    })
}

HAPI double mllm_cuda_llama_last_forward_ms(void *h) { return ((HostModel *)h)->last_ms; }
HAPI long long mllm_cuda_kernel_launches() {
    try { return backend()->kernel_launches; } catch (...) { return -1; }
}
HAPI int mllm_cuda_pool_stats(size_t *bytes, size_t *hits, size_t *misses) {
    GUARD({
        CUDABackend *b = backend();
        *bytes = b->bytes_allocated; *hits = b->pool_hits; *misses = b->pool_misses;
    })
}
HAPI int mllm_cuda_sync() { GUARD({ backend()->sync(); }) }

// ---- per-op access for the per-op parity tests (test/cpu/CPUTest.hpp pattern: construct an Op
// from the backend, reshape -> setUp -> load -> execute, compare tensors).
// Runs ONE op given OpParam as parallel key/value arrays; inputs are host fp32 buffers with
// shapes [n_in][4] (b,h,s,d); weights come from the .mllm at `path` (may be empty).
HAPI int mllm_cuda_run_op(const char **keys, const float *vals, int n_params, const char *name, const char *path,
                          const float **inputs, const int *in_shapes, int n_in, float *out_host, int *out_shape) {
    GUARD({
        CUDABackend *bn = backend();
        OpParam p;
        int n_out = 1, out_index = 0; // test-harness keys, not op parameters: ops with several outputs (split) return the chosen one
        for (int i = 0; i < n_params; i++) {
            if (string(keys[i]) == "__n_out") n_out = (int)vals[i];
            else if (string(keys[i]) == "__out_index") out_index = (int)vals[i];
            else p[keys[i]] = vals[i];
        }
        std::unique_ptr<Op> op(bn->opCreate(p, name));
        ParamLoader loader(path ? path : "");
        op->load(loader);
        vector<Tensor> in;
        for (int i = 0; i < n_in; i++) {
            const int *s = in_shapes + 4 * i;
            Tensor t(s[0], s[1], s[2], s[3], bn);
            t.setName(string("in") + std::to_string(i));
            t.copyFromHost(inputs[i]);
            in.push_back(t);
        }
        vector<string> out_names;
        for (int i = 0; i < n_out; i++) out_names.push_back(string("out-") + name + (n_out > 1 ? "-" + std::to_string(i) : ""));
        vector<Tensor> out = bn->runOp(op.get(), in, out_names, false);
        Tensor &o = out.at(out_index);
        out_shape[0] = o.batch(); out_shape[1] = o.head(); out_shape[2] = o.sequence(); out_shape[3] = o.dimension();
        if (o.dtype() != MLLM_TYPE_F32 || !o.contiguous()) throw std::runtime_error("run_op: output must be dense fp32");
        o.copyToHost(out_host);
    })
}

// Tensor::half() / Tensor::fp32() through Backend::convert_fp_data: x [rows][cols] fp32 -> fp16 bits -> fp32 again (test access)
HAPI int mllm_cuda_convert_fp_roundtrip(const float *x, int rows, int cols, uint16_t *half_out, float *back_out) {
    GUARD({
        CUDABackend *bn = backend();
        Tensor a(1, 1, rows, cols, bn), h(bn), b(bn);
        a.copyFromHost(x);
        h.reshape(1, 1, rows, cols); h.setDtype(MLLM_TYPE_F16); h.alloc();
        b.reshape(1, 1, rows, cols); b.setDtype(MLLM_TYPE_F32); b.alloc();
        bn->convert_fp_data(&a, &h);
        bn->convert_fp_data(&h, &b);
        h.copyToHost(half_out);
        b.copyToHost(back_out);
    })
}

HAPI void *mllm_cuda_stream() {
    try { return backend()->stream(); } catch (...) { return nullptr; }
}

// One request with inputs already resident in HBM (what bench.py's `value` times): ids are uploaded
// BEFORE the timed region; prefill of S tokens, then n_decode single-token steps whose input is the
// previous step's on-device argmax (mllmcu_argmax_ids) -- no host<->device traffic inside the region.
// Device time (CUDA events on the backend stream) of the two phases is returned in ms.
// tokens_out (host, optional): the B*n_decode greedy tokens, copied back after the timed region.
HAPI int mllm_cuda_llama_request_device(void *hv, const float *ids_host, int B, int S, int n_decode, float *ms_prefill,
                                        float *ms_decode, int *tokens_out) {
    GUARD({
        auto *h = (HostModel *)hv;
        CUDABackend *bn = h->model->backend();
        const int V = h->model->config().vocab;
        Tensor ids(B, 1, S, 1, bn);
        ids.setName("input_ids");
        float *stage = (float *)h->pinned(sizeof(float) * (size_t)B * S); // the copy engine reads pinned memory directly
        memcpy(stage, ids_host, sizeof(float) * (size_t)B * S);
        ids.copyFromHost(stage);
        Tensor toks(1, 1, n_decode > 0 ? n_decode : 1, B, bn); // fp32 ids, row per step
        const int len0 = h->model->cache_seq_len();            // a request may continue a cache (multi-turn, Module::generate without clear)
        void *e0, *e1, *e2;
        MLLMCU_CHECK(mllmcu_event_create(&e0));
        MLLMCU_CHECK(mllmcu_event_create(&e1));
        MLLMCU_CHECK(mllmcu_event_create(&e2));
        bn->sync();
        MLLMCU_CHECK(mllmcu_event_record(e0, bn->stream()));
        Tensor lg = h->model->forward(ids);
        if (n_decode > 0) MLLMCU_CHECK(mllmcu_argmax_ids(lg.devPtr<float>(), B, V, toks.devPtr<float>(), bn->stream()));
        MLLMCU_CHECK(mllmcu_event_record(e1, bn->stream()));
        std::vector<char> fused((size_t)(n_decode > 0 ? n_decode : 1), 0);
        const float *cur_ids = toks.devPtr<float>(); // ids consumed by the next step
        for (int step = 1; step < n_decode; step++) {
            if (h->model->decode_plan_usable(B)) { // fused step + on-device pick: one graph launch per token
                h->model->decode_step(cur_ids, B, true);
                cur_ids = h->model->decode_ids();
                fused[step] = 1;
                continue;
            }
            Tensor cur(B, 1, 1, 1, bn);
            cur.setName("input_ids");
            MLLMCU_CHECK(mllmcu_memcpy_d2d(cur.rawDevPtr(), cur_ids, sizeof(float) * B, bn->stream()));
            Tensor l2 = h->model->forward(cur);
            MLLMCU_CHECK(mllmcu_argmax_ids(l2.devPtr<float>(), B, V, toks.devPtr<float>() + (size_t)step * B, bn->stream()));
            cur_ids = toks.devPtr<float>() + (size_t)step * B;
        }
        MLLMCU_CHECK(mllmcu_event_record(e2, bn->stream()));
        MLLMCU_CHECK(mllmcu_event_sync(e2));
        MLLMCU_CHECK(mllmcu_event_elapsed_ms(e0, e1, ms_prefill));
        MLLMCU_CHECK(mllmcu_event_elapsed_ms(e1, e2, ms_decode));
        mllmcu_event_destroy(e0); mllmcu_event_destroy(e1); mllmcu_event_destroy(e2);
        if (tokens_out && n_decode > 0) {
            std::vector<float> t((size_t)n_decode * B);
            toks.copyToHost(t.data());
            std::vector<int> log;
            const int *log_dev = h->model->decode_tok_log();
            if (log_dev) {
                log.resize((size_t)(h->model->config().cache_limit + 1) * B);
                MLLMCU_CHECK(mllmcu_memcpy_d2h(log.data(), log_dev, log.size() * sizeof(int), bn->stream()));
                bn->sync();
            }
            for (int s = 0; s < n_decode; s++)
                for (int b = 0; b < B; b++) // step s consumed position len0+S+s-1; its pick is logged under that position
                    tokens_out[(size_t)b * n_decode + s] = fused[s] ? log[(size_t)(len0 + S + s - 1) * B + b] : (int)t[(size_t)s * B + b];
        }
    })
}

// 1 when the last forward() of this model ran as the fused (graph) decode step, 0 when it ran op by op
HAPI int mllm_cuda_llama_last_step_fused(void *hv) { return ((HostModel *)hv)->model->last_step_fused() ? 1 : 0; }
HAPI int mllm_cuda_llama_set_eager_decode(void *hv, int on) { GUARD({ ((HostModel *)hv)->model->set_eager_decode(on != 0); }) }
HAPI int mllm_cuda_llama_decode_mode(void *hv) { return ((HostModel *)hv)->model->decode_mode(); }
HAPI int mllm_cuda_llama_set_step_kernel(void *hv, int on) { GUARD({ ((HostModel *)hv)->model->set_step_kernel(on != 0); }) }
