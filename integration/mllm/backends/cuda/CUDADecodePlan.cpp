// mllm/backends/cuda/CUDADecodePlan.cpp -- the fused single-token step behind an UNCHANGED model header.
//
// The reference executes a decode token as ~25 runOp calls per transformer block (Layer::run -> Backend::runOp,
// Layer.hpp:116-222); on a B200 that is ~770 launches and ~6 ms of host dispatch per token for the 7B.  The CUDA backend
// therefore RECORDS the op sequence of the first single-token forward it sees (runOp appends every executed op), recognises the
// LLaMA-family decoder pattern
//     EMBEDDING, N x [RMSNORM, LINEAR q, LINEAR k, LINEAR v, ROPE q, ROPE k, KVCACHE k, KVCACHE v, F_FA2, LINEAR o, F_TTADD,
//                     RMSNORM, LINEAR gate, SILU, LINEAR up, F_TTMUL, LINEAR down, F_TTADD], RMSNORM, LINEAR lm_head
// (F_VIEW is metadata and ignored; this is LLaMAModel, TinyLLaMAModel and QWenForCausalLM as their headers issue them), reads
// each op's device state through CUDAOpBase::fuseInfo, and from then on answers single-token forwards with the fused launches of
// the decode step -- five per block, captured once in a CUDA graph with the position held in device memory -- without calling
// Module::Forward.  The ops stay the owners of state: after every replay the RoPE counters and KV-cache lengths are advanced, so
// op-by-op forwards (the next prefill, a chunk) and fused steps interleave freely.  Anything that does not match the pattern or
// the fused kernels' shapes simply keeps running op by op.
#include "CUDABackend.hpp"

#include "Module.hpp"
#include "Tensor.hpp"
#include "backends/cpu/op/CPUKVCache.hpp"
#include "mllmcu.h"

#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <iostream>

namespace mllm {

struct CUDABackend::DecodePlan {
    struct L {
        CUDAFuseInfo n1, n2, q, k, v, o, gate, up, down, rope;
        CUDAOpBase *q_rope = nullptr, *k_rope = nullptr;
        CPUKVCache *kc = nullptr, *vc = nullptr;
        void *kptr = nullptr, *vptr = nullptr; // cache addresses the captured graph holds
    };
    Module *module = nullptr;
    int B = 0, H = 0, Hq = 0, Hkv = 0, D = 0, F = 0, V = 0, limit = 0;
    long long kv_bstride = 0;
    std::vector<L> layers;
    CUDAFuseInfo embd, norm, head;
    Tensor x, x2, qkv, ao, g, logits, ids, ctl, attn_ws;
    Tensor t1, t2, gu, lin_ws; // k-quant models: normalised row, Linear output, up projection, the Linears' Q8_K workspace
    size_t attn_ws_bytes = 0, lin_ws_bytes = 0;
    bool kquant = false;       // Linears are Q4_K / Q6_K: op-level kernels on fixed buffers inside the captured graph
    bool q4k_chain = true;     // ... every one of them Q4_K: the nine-launch chain of kquants.cu (prep + multi-matrix GEMV, PDL)
    bool mrope = false;        // Qwen2-VL: MultimodalRoPE ops; the rotary position comes with the call (position ids), not from the cache length
    int rope_delta = 0;        // rotary position - cache length of the request being decoded (constant over its decode steps)
    Tensor msin, mcos;         // mrope: rotate-half tables [limit][D] for text positions
    int launches = 0;
    int dev_pos = -2, ctl_host = 0;
    bool warmed = false, use_graph = true, use_pdl = true, unsupported = false;
    void *graph = nullptr;
    float *stage_ids = nullptr; // pinned
    ~DecodePlan() {
        if (graph) mllmcu_graph_destroy(graph);
        if (stage_ids) mllmcu_host_free(stage_ids);
    }
};

static bool env_on(const char *k) {
    const char *e = getenv(k);
    return e && e[0] && e[0] != '0';
}

struct DecodePlanBuilder {
    // builds a plan from the recorded ops, or returns nullptr when the sequence is not the pattern above
    static std::shared_ptr<CUDABackend::DecodePlan> build(CUDABackend *bn, Module *module, const std::vector<Op *> &rec, int B) {
        std::vector<Op *> ops;
        for (Op *o : rec)
            if (o->type() != F_VIEW) ops.push_back(o);
        // LLaVAModel::Forward ends with clip({}, {}, {-1}, {}) (modeling_llava.hpp:135): on a one-token forward that is the identity
        if (!ops.empty() && ops.back()->type() == F_CLIP) ops.pop_back();
        static const OpType block[18] = {RMSNORM, LINEAR, LINEAR, LINEAR, ROPE, ROPE, KVCACHE, KVCACHE, F_FA2, LINEAR, F_TTADD,
                                         RMSNORM, LINEAR, SILU, LINEAR, F_TTMUL, LINEAR, F_TTADD};
        // QWen2Decoder of Qwen2-VL (modeling_qwen2_vl.hpp:225-262,310-330) is the same block with MultimodalRoPE in the two rotary slots
        const bool mrope = ops.size() > 6 && ops[5]->type() == MULTIMODALROPE;
        if (ops.size() < 3 + 18 || (ops.size() - 3) % 18 != 0) return nullptr;
        const int nl = (int)(ops.size() - 3) / 18;
        if (ops[0]->type() != EMBEDDING || ops[ops.size() - 2]->type() != RMSNORM || ops[ops.size() - 1]->type() != LINEAR) return nullptr;
        for (int l = 0; l < nl; l++)
            for (int j = 0; j < 18; j++)
                if (ops[1 + l * 18 + j]->type() != ((mrope && block[j] == ROPE) ? MULTIMODALROPE : block[j])) return nullptr;
        auto p = std::make_shared<CUDABackend::DecodePlan>();
        auto fi = [](Op *o, CUDAFuseInfo &f) {
            auto *c = dynamic_cast<CUDAOpBase *>(o);
            return c && c->fuseInfo(f);
        };
        if (!fi(ops[0], p->embd) || !fi(ops[ops.size() - 2], p->norm) || !fi(ops[ops.size() - 1], p->head)) return nullptr;
        p->layers.resize(nl);
        p->mrope = mrope;
        for (int l = 0; l < nl; l++) {
            Op **o = &ops[1 + l * 18];
            auto &L = p->layers[l];
            if (!(fi(o[0], L.n1) && fi(o[1], L.q) && fi(o[2], L.k) && fi(o[3], L.v) && fi(o[4], L.rope) && fi(o[9], L.o) && fi(o[11], L.n2) &&
                  fi(o[12], L.gate) && fi(o[14], L.up) && fi(o[16], L.down)))
                return nullptr;
            L.q_rope = dynamic_cast<CUDAOpBase *>(o[4]);
            L.k_rope = dynamic_cast<CUDAOpBase *>(o[5]);
            L.kc = dynamic_cast<CPUKVCache *>(o[6]);
            L.vc = dynamic_cast<CPUKVCache *>(o[7]);
            if (!L.q_rope || !L.k_rope || !L.kc || !L.vc) return nullptr;
            L.kptr = L.kc->cache_->rawHostPtr();
            L.vptr = L.vc->cache_->rawHostPtr();
            for (const CUDAFuseInfo *f : {&L.q, &L.k, &L.v, &L.o, &L.gate, &L.up, &L.down}) {
                if (f->wtype == MLLM_TYPE_Q4_K || f->wtype == MLLM_TYPE_Q6_K) p->kquant = true;
                else if (f->wtype != MLLM_TYPE_Q4_0) return nullptr;
                if (f->wtype != MLLM_TYPE_Q4_K) p->q4k_chain = false;
            }
            if (L.n1.unit_offset || L.n2.unit_offset) return nullptr;
        }
        const auto &L0 = p->layers[0];
        Tensor &kc = *L0.kc->cache_;
        if (kc.dtype() != MLLM_TYPE_F16 || kc.ctype() != BSHD || kc.batch() != B) return nullptr;
        p->module = module;
        p->B = B;
        p->H = L0.q.in;
        p->D = kc.dimension();
        p->Hkv = kc.head();
        p->Hq = L0.q.out / p->D;
        p->F = L0.gate.out;
        p->V = p->head.out;
        p->limit = kc.sequence();
        p->kv_bstride = (long long)kc.sequence() * kc.head() * kc.dimension();
        if (p->head.wtype == MLLM_TYPE_Q4_K || p->head.wtype == MLLM_TYPE_Q6_K) p->kquant = true;
        else if (p->head.wtype != MLLM_TYPE_Q4_0) return nullptr;
        if (p->head.wtype != MLLM_TYPE_Q4_K || getenv("MLLM_CUDA_Q4K_CHAIN_OFF")) p->q4k_chain = false;
        if (p->embd.wtype != MLLM_TYPE_Q4_0 || p->norm.unit_offset) return nullptr;
        if (p->kquant && (p->H % 256 || p->F % 256 || (p->Hq * p->D) % 256)) return nullptr; // Q8_K super-blocks
        if (p->H % 128 || p->F % 128 || (p->D != 32 && p->D != 64 && p->D != 128) || L0.rope.rot_dim != p->D || B < 1 || B > 8) return nullptr;
        if (mrope) { // one rotate-half table for text positions 0 .. limit - 1, as CUDARoPE / CPURoPE build it (theta in double -> float, libm)
            if (B != 1 || p->kquant) return nullptr;
            const int D = p->D, half = D / 2;
            std::vector<float> theta(half), sn((size_t)p->limit * D), cs((size_t)p->limit * D);
            for (int i = 0; i < half; i++) theta[i] = (float)(1.0 / pow((double)L0.rope.rope_theta, 2.0 * i / D));
            for (int pos = 0; pos < p->limit; pos++)
                for (int i = 0; i < half; i++) {
                    const float value = theta[i] * (float)pos; // CPUMultimodalRoPE.cpp: theta * position id (a float)
                    sn[(size_t)pos * D + i] = sn[(size_t)pos * D + i + half] = sinf(value);
                    cs[(size_t)pos * D + i] = cs[(size_t)pos * D + i + half] = cosf(value);
                }
            auto up = [&](Tensor &t, const std::vector<float> &src) {
                t.setBackend(bn);
                t.reshape(1, 1, p->limit, D);
                t.setDtype(MLLM_TYPE_F32);
                bn->memory()->device_only = true;
                t.alloc();
                bn->memory()->device_only = false;
                MLLMCU_CHECK(mllmcu_memcpy_h2d(t.rawHostPtr(), src.data(), src.size() * sizeof(float), bn->stream()));
                bn->sync();
            };
            up(p->msin, sn);
            up(p->mcos, cs);
        }
        for (auto &L : p->layers) {
            Tensor &a = *L.kc->cache_, &b = *L.vc->cache_;
            if (a.batch() != B || b.batch() != B || a.sequence() != p->limit || b.sequence() != p->limit || a.head() != p->Hkv || a.dimension() != p->D ||
                L.rope.rot_dim != p->D || L.rope.sin_t != L0.rope.sin_t || L.rope.rope_theta != L0.rope.rope_theta)
                return nullptr;
        }
        if (mrope)
            for (auto &L : p->layers) { L.rope.sin_t = p->msin.hostPtr<float>(); L.rope.cos_t = p->mcos.hostPtr<float>(); L.rope.pos_max = p->limit; }
        auto mk = [&](Tensor &t, int rows, int cols) {
            t.setBackend(bn);
            t.reshape(1, 1, rows, cols);
            t.setDtype(MLLM_TYPE_F32);
            bn->memory()->device_only = true; // plan buffers never leave the device
            t.alloc();
            bn->memory()->device_only = false;
        };
        const int QKV = (p->Hq + 2 * p->Hkv) * p->D;
        mk(p->x, B, p->H); mk(p->x2, B, p->H); mk(p->qkv, B, QKV); mk(p->ao, B, p->Hq * p->D); mk(p->g, B, p->F); mk(p->logits, B, p->V);
        mk(p->ids, 1, B); mk(p->ctl, 1, 4);
        if (p->kquant) {
            const int widest = std::max(std::max(p->H, p->F), QKV);
            mk(p->t1, B, widest); mk(p->t2, B, std::max(p->H, p->V)); mk(p->gu, B, p->F);
            p->lin_ws_bytes = std::max(mllmcu_linear_workspace_bytes(B, widest), (size_t)B * (widest / 256) * 304 + 256); // 304: the chain's re-pitched Q8_K blocks
            mk(p->lin_ws, 1, (int)((p->lin_ws_bytes + 3) / 4));
        }
        p->attn_ws_bytes = mllmcu_decode_attn_workspace_bytes(B, p->Hq, p->D);
        mk(p->attn_ws, 1, (int)((p->attn_ws_bytes + 3) / 4));
        MLLMCU_CHECK(mllmcu_memset(p->attn_ws.rawHostPtr(), 0, p->attn_ws.cntSize(), bn->stream()));
        MLLMCU_CHECK(mllmcu_memset(p->ctl.rawHostPtr(), 0, 16, bn->stream()));
        void *hp = nullptr;
        MLLMCU_CHECK(mllmcu_host_alloc(&hp, sizeof(float) * 8));
        p->stage_ids = (float *)hp;
        p->use_graph = !env_on("MLLM_CUDA_NO_GRAPH");
        p->use_pdl = !env_on("MLLM_CUDA_NO_PDL");
        return p;
    }

    static bool issue_kquant(CUDABackend *bn, CUDABackend::DecodePlan &p);
    // the kernel sequence of one token (five launches per block); false when a fused kernel refuses the shape
    static bool issue(CUDABackend *bn, CUDABackend::DecodePlan &p, bool pdl_on) {
        void *st = bn->stream();
        const int B = p.B, H = p.H, D = p.D, Hq = p.Hq, Hkv = p.Hkv, F = p.F, QKV = (Hq + 2 * Hkv) * D, pdl = pdl_on ? 1 : 0;
        int *ctl = (int *)p.ctl.rawHostPtr();
        float *x = p.x.hostPtr<float>(), *x2 = p.x2.hostPtr<float>(), *qkv = p.qkv.hostPtr<float>(), *ao = p.ao.hostPtr<float>(), *g = p.g.hostPtr<float>();
        auto gemv = [&](std::initializer_list<const CUDAFuseInfo *> ws, const float *in, int K, const CUDAFuseInfo *norm, int epi, const float *res,
                        float *out, int ldy, bool late_trigger = false) -> bool {
            mllmcu_gemv_desc d = {};
            d.wtype = MLLM_TYPE_Q4_0;
            for (const CUDAFuseInfo *f : ws) { d.W[d.nW] = f->weight; d.N[d.nW] = f->out; d.bias[d.nW] = f->bias; d.nW++; }
            d.x = in; d.ldx = K; d.M = B; d.K = K;
            if (norm) { d.norm_w = norm->norm_w; d.eps = norm->eps; d.norm_reps = 1; d.norm_rep_stride = K; }
            d.epilogue = epi; d.residual = res; d.ldr = ldy; d.y = out; d.ldy = ldy; d.pdl = (pdl && late_trigger) ? 2 : pdl;
            d.x_reps = 1; d.y_reps = 1;
            const int rc = mllmcu_gemv_fused_ex(&d, st);
            if (rc == -2) return false;
            MLLMCU_CHECK(rc);
            return true;
        };
        MLLMCU_CHECK(mllmcu_decode_embed(p.embd.weight, p.ids.hostPtr<float>(), x, B, H, ctl, st));
        if (p.kquant) return issue_kquant(bn, p);
        p.launches = 2 + 5 * (int)p.layers.size();
        for (size_t li = 0; li < p.layers.size(); li++) {
            auto &l = p.layers[li];
            // layer 0's projection releases its dependents only after its own wait: every later kernel then starts with the embedding
            // kernel (which bumps ctl[0]) complete and may read the position and its K/V rows ahead of its own wait
            if (!gemv({&l.q, &l.k, &l.v}, x, H, &l.n1, 0, nullptr, qkv, QKV, li == 0)) return false;
            // the kernel rotates with table row ctl[0] (the cache length); a Qwen2-VL text token sits at rotary position ctl[0] + rope_delta,
            // so the table base is moved by that many rows (rows in use stay inside [0, limit): checked before the launch)
            const long long shift = (long long)p.rope_delta * l.rope.rot_dim;
            MLLMCU_CHECK(mllmcu_decode_attn_rope(l.rope.rope_type, qkv, QKV, l.kptr, l.vptr, ao, Hq * D, ctl, B, Hq,
                                                 Hkv, D, l.rope.sin_t + shift, l.rope.cos_t + shift, l.rope.rot_dim, p.kv_bstride, p.limit, p.attn_ws.rawHostPtr(),
                                                 p.attn_ws_bytes, pdl ? 2 : 0, st));
            if (!gemv({&l.o}, ao, Hq * D, nullptr, 1, x, x2, H)) return false;
            if (!gemv({&l.gate, &l.up}, x2, H, &l.n2, 2, nullptr, g, F)) return false;
            if (!gemv({&l.down}, g, F, nullptr, 1, x2, x, H)) return false;
        }
        return gemv({&p.head}, x, H, &p.norm, 0, nullptr, p.logits.hostPtr<float>(), p.V);
    }
};

// k-quant models (the LLaVA body is Q4_K).  All Linears Q4_K: nine launches per block -- the op in front of each Linear (RMSNorm, SiLU * up) folded
// into the block_q8_K quantiser, q | k | v and gate | up as one GEMV launch each, the residual add in the epilogue of o / down, every kernel a
// programmatic dependent of the one before it (kquants.cu, mllmcu_prep_q8_K / mllmcu_gemv_q4k; same quantised bytes and integers as the op-by-op sequence, another fp32 summation order).
// Mixed Q4_K / Q6_K / Q4_0 bodies: the same step from the op-level kernels -- RMSNorm, Linear, the fused RoPE + append + attention kernel, add,
// SiLU * mul -- on the plan's fixed buffers.  Either way the sequence is captured once and replayed without host dispatch.
bool DecodePlanBuilder::issue_kquant(CUDABackend *bn, CUDABackend::DecodePlan &p) {
    void *st = bn->stream();
    const int B = p.B, H = p.H, D = p.D, Hq = p.Hq, Hkv = p.Hkv, F = p.F, QKV = (Hq + 2 * Hkv) * D;
    int *ctl = (int *)p.ctl.rawHostPtr();
    float *x = p.x.hostPtr<float>(), *x2 = p.x2.hostPtr<float>(), *qkv = p.qkv.hostPtr<float>(), *ao = p.ao.hostPtr<float>(), *g = p.g.hostPtr<float>();
    float *t1 = p.t1.hostPtr<float>(), *t2 = p.t2.hostPtr<float>(), *gu = p.gu.hostPtr<float>();
    void *ws = p.lin_ws.rawHostPtr();
    int n = 1;
    if (p.q4k_chain) {
        auto prep = [&](int mode, const float *in, const float *aux, float eps, int K) {
            MLLMCU_CHECK(mllmcu_prep_q8_K(mode, in, aux, eps, 0, ws, B, K, K, 3, st)); // 3: PDL + 304-byte blocks
            n++;
        };
        auto gemv = [&](std::initializer_list<const CUDAFuseInfo *> fs, std::initializer_list<float *> outs, int ldy, int K, const float *res) {
            const void *W[3] = {};
            const float *bias[3] = {};
            float *y[3] = {};
            int N[3] = {}, k = 0;
            for (const CUDAFuseInfo *f : fs) { W[k] = f->weight; bias[k] = f->bias; N[k] = f->out; k++; }
            k = 0;
            for (float *o : outs) y[k++] = o;
            MLLMCU_CHECK(mllmcu_gemv_q4k(k, W, N, bias, y, ldy, ws, B, K, res, ldy, 3, st));
            n++;
        };
        for (auto &l : p.layers) {
            prep(1, x, l.n1.norm_w, l.n1.eps, H);
            gemv({&l.q, &l.k, &l.v}, {qkv, qkv + (size_t)Hq * D, qkv + (size_t)(Hq + Hkv) * D}, QKV, H, nullptr);
            MLLMCU_CHECK(mllmcu_decode_attn_rope(l.rope.rope_type, qkv, QKV, l.kptr, l.vptr, ao, Hq * D, ctl, B, Hq, Hkv, D, l.rope.sin_t, l.rope.cos_t,
                                                 l.rope.rot_dim, p.kv_bstride, p.limit, p.attn_ws.rawHostPtr(), p.attn_ws_bytes, 1, st));
            n++;
            prep(0, ao, nullptr, 0.f, Hq * D);
            gemv({&l.o}, {x2}, H, Hq * D, x);
            prep(1, x2, l.n2.norm_w, l.n2.eps, H);
            gemv({&l.gate, &l.up}, {g, gu}, F, H, nullptr);
            prep(2, g, gu, 0.f, F);
            gemv({&l.down}, {x}, H, F, x2);
        }
        prep(1, x, p.norm.norm_w, p.norm.eps, H);
        gemv({&p.head}, {p.logits.hostPtr<float>()}, p.V, H, nullptr);
        p.launches = n;
        return true;
    }
    auto lin = [&](const CUDAFuseInfo &f, const float *in, int K, float *out, int ldy, int reuse) {
        MLLMCU_CHECK(mllmcu_linear_ex(f.wtype, f.weight, in, f.bias, out, B, f.out, K, K, ldy, 0, ws, p.lin_ws_bytes, reuse, st));
        n += 2 - reuse;
    };
    auto kq = [](const CUDAFuseInfo &f) { return f.wtype == MLLM_TYPE_Q4_K || f.wtype == MLLM_TYPE_Q6_K; };
    for (auto &l : p.layers) {
        MLLMCU_CHECK(mllmcu_rmsnorm(x, l.n1.norm_w, t1, B, H, l.n1.eps, 0, st));
        lin(l.q, t1, H, qkv, QKV, 0);
        lin(l.k, t1, H, qkv + (size_t)Hq * D, QKV, kq(l.q) && kq(l.k)); // q / k / v read one block_q8_K image of the normalised row
        lin(l.v, t1, H, qkv + (size_t)(Hq + Hkv) * D, QKV, kq(l.q) && kq(l.k) && kq(l.v));
        MLLMCU_CHECK(mllmcu_decode_attn_rope(l.rope.rope_type, qkv, QKV, l.kptr, l.vptr, ao, Hq * D, ctl, B, Hq, Hkv, D, l.rope.sin_t, l.rope.cos_t,
                                             l.rope.rot_dim, p.kv_bstride, p.limit, p.attn_ws.rawHostPtr(), p.attn_ws_bytes, 0, st));
        lin(l.o, ao, Hq * D, t2, H, 0);
        MLLMCU_CHECK(mllmcu_add(x, t2, x2, (size_t)B * H, st));
        MLLMCU_CHECK(mllmcu_rmsnorm(x2, l.n2.norm_w, t1, B, H, l.n2.eps, 0, st));
        lin(l.gate, t1, H, g, F, 0);
        lin(l.up, t1, H, gu, F, kq(l.gate) && kq(l.up));
        MLLMCU_CHECK(mllmcu_silu_mul(g, gu, g, (size_t)B * F, st));
        lin(l.down, g, F, t2, H, 0);
        MLLMCU_CHECK(mllmcu_add(x2, t2, x, (size_t)B * H, st));
        n += 6;
    }
    MLLMCU_CHECK(mllmcu_rmsnorm(x, p.norm.norm_w, t1, B, H, p.norm.eps, 0, st));
    lin(p.head, t1, H, p.logits.hostPtr<float>(), p.V, 0);
    p.launches = n + 1;
    return true;
}

// Called by runForward for every outermost forward.  Returns true when `out` holds the logits of a fused step.
bool CUDABackend::try_decode_plan(Module *module, std::vector<Tensor> &inputs, std::vector<std::any> &args, std::vector<Tensor> &out) {
    if (!fused_decode || inputs.empty()) return false;
    // LLaVAModel's decode call is {ids, image emptied by chatPostProcessing} (demo_llava.cpp:57-61); Qwen2VLModel's is {ids, pixels and
    // grid emptied, position ids [3,1,1,1]} (demo_qwen2_vl.cpp:47-64, Qwen2VLModel::get_position_ids)
    if (inputs.size() > 4) return false;
    for (size_t i = 1; i < inputs.size() && i < 3; i++)
        if (inputs[i].count() != 0) return false;
    if (inputs.size() == 4 && (inputs[3].batch() != 3 || inputs[3].count() != 3 || inputs[3].dtype() != MLLM_TYPE_F32)) return false;
    Tensor &in = inputs[0];
    const int B = in.batch();
    if (in.sequence() != 1 || in.head() != 1 || in.dimension() != 1 || in.dtype() != MLLM_TYPE_F32 || B < 1 || B > 8) return false;
    if (plan_ && (plan_->module != module || plan_->B != B)) { sync(); plan_.reset(); }
    if (plan_ && plan_->unsupported) return false;
    if (!plan_) {
        // record this forward: it runs op by op (and is the result of this call); the plan, if the sequence matches, serves the next ones
        recording_ = true;
        rec_ops_.clear();
        std::vector<Tensor> res;
        try { res = module->Forward(inputs, args); } catch (...) { recording_ = false; throw; }
        recording_ = false;
        plan_ = DecodePlanBuilder::build(this, module, rec_ops_, B);
        if (!plan_) { plan_ = std::make_shared<DecodePlan>(); plan_->module = module; plan_->B = B; plan_->unsupported = true; }
        rec_ops_.clear();
        out = res;
        return true;
    }
    DecodePlan &p = *plan_;
    // per-step state: every RoPE counter and cache length must agree, and there must be room
    const int len = p.layers[0].kc->getCacheSeqLen();
    for (auto &l : p.layers) {
        if (l.kc->getCacheSeqLen() != len || l.vc->getCacheSeqLen() != len) return false;
        if (l.kc->cache_->rawHostPtr() != l.kptr || l.vc->cache_->rawHostPtr() != l.vptr) { // the cache was re-made (batch change): record again
            sync();
            plan_.reset();
            return false;
        }
    }
    if (len + 1 > p.limit) { // CPUKVCache.cpp:121-126
        std::cerr << "\n[ERROR]: Current tokens exceed cache limit: " << len + 1 << ">" << p.limit << ";"
                  << "\n         Please set args `--limits` >" << p.limit << std::endl;
        throw std::runtime_error("[mllm-cuda] KV cache limit exceeded");
    }
    if (p.mrope != (inputs.size() == 4)) return false;
    if (p.mrope) { // a text token: the three axes carry one position; its distance to the cache length is fixed for the request
        sync();
        const float pt = inputs[3].dataAt<float>(0, 0, 0, 0), ph = inputs[3].dataAt<float>(1, 0, 0, 0), pw = inputs[3].dataAt<float>(2, 0, 0, 0);
        if (pt != ph || pt != pw || pt < 0 || pt >= (float)p.limit || pt != (float)(int)pt) return false;
        const int delta = (int)pt - len;
        if (delta != p.rope_delta) { // another request: the captured launches hold the old table base
            if (p.graph) { mllmcu_graph_destroy(p.graph); p.graph = nullptr; }
            p.rope_delta = delta;
        }
    } else if (len + 1 > p.layers[0].rope.pos_max) throw std::runtime_error("[mllm-cuda] RoPE: position exceeds max_position_embeddings");
    dropQuantTag();
    // token ids: the demos build the input on the CPU backend (host memory); a CUDA-backend tensor is copied device to device
    if (in.backend()->type() == MLLM_CPU) {
        for (int b = 0; b < B; b++) p.stage_ids[b] = in.dataAt<float>(b, 0, 0, 0);
        MLLMCU_CHECK(mllmcu_memcpy_h2d(p.ids.rawHostPtr(), p.stage_ids, sizeof(float) * B, stream_));
    } else {
        MLLMCU_CHECK(mllmcu_memcpy_d2d(p.ids.rawHostPtr(), in.ptrAt<float>(0, 0, 0, 0), sizeof(float) * B, stream_));
    }
    if (p.dev_pos != len - 1) { // the embedding kernel pre-increments ctl[0]
        p.ctl_host = len - 1;
        MLLMCU_CHECK(mllmcu_memcpy_h2d(p.ctl.rawHostPtr(), &p.ctl_host, sizeof(int), stream_));
        p.dev_pos = len - 1;
    }
    bool ran = false;
    if (p.use_graph && !p.graph && p.warmed) { // second fused step: every kernel attribute is configured, capture the sequence
        for (int attempt = 0; attempt < 2 && !p.graph; attempt++) {
            MLLMCU_CHECK(mllmcu_graph_begin_capture(stream_));
            bool ok = false;
            try { ok = DecodePlanBuilder::issue(this, p, p.use_pdl); } catch (...) { mllmcu_graph_abort_capture(stream_); throw; }
            void *ge = nullptr;
            if (!ok) { mllmcu_graph_abort_capture(stream_); break; }
            if (mllmcu_graph_end_capture(stream_, &ge) == 0) p.graph = ge;
            else if (p.use_pdl) p.use_pdl = false;
            else { p.use_graph = false; break; }
        }
    }
    if (p.graph) {
        MLLMCU_CHECK(mllmcu_graph_launch(p.graph, stream_));
        ran = true;
    } else {
        ran = DecodePlanBuilder::issue(this, p, p.use_pdl);
        p.warmed = true;
    }
    if (!ran) { // a kernel refused the shape mid-step (nothing stateful ran before it): this model stays op by op
        p.unsupported = true;
        p.dev_pos = -2;
        return false;
    }
    p.dev_pos = len;
    // the ops own the state: advance RoPE positions and cache lengths by one token
    static Tensor one_token(1, 1, 1, 1, Backend::global_backends[MLLM_CPU].get(), false);
    auto one = std::shared_ptr<Tensor>(&one_token, [](Tensor *) {});
    for (auto &l : p.layers) {
        l.q_rope->advance(1);
        l.k_rope->advance(1);
        l.kc->execute({one}, {}); // n_rep == 1: only cache_seq_len_ += inputs[0]->sequence() (CPUKVCache.cpp:148-151)
        l.vc->execute({one}, {});
    }
    kernel_launches += p.launches;
    last_forward_fused = 1;
    // hand out a fresh tensor (callers may migrate it with .cpu(), which would take the plan's buffer with it)
    Tensor lg(B, 1, 1, p.V, this, true);
    lg.setName("out-lm_head-fused");
    lg.setModule(module);
    MLLMCU_CHECK(mllmcu_memcpy_d2d(lg.rawHostPtr(), p.logits.rawHostPtr(), sizeof(float) * (size_t)B * p.V, stream_));
    out = {lg};
    return true;
}

} // namespace mllm
