// ref_ops_shim.cpp -- C-ABI over the REFERENCE's own CPU backend (TEST INFRASTRUCTURE ONLY).
//
// This file contains no arithmetic of its own.  It is linked with the reference's CPU backend compiled from the sources
// where they lie under /root/reference (oracle/cpuref.mk) and exposes, for the Python tests and golden-vector scripts,
//   (i)  any reference Op class, created the way test/cpu/CPUTest.hpp:12-53 does (CPUBackend::opCreate with an OpParam
//        map, Op::load from a ParamLoader over a .mllm file, reshape -> setUp -> execute), and
//   (ii) the reference's model graphs run on MLLM_CPU exactly as examples/demo_llama.cpp / demo_tinyllama.cpp /
//        demo_qwen.cpp run them: LLaMAModel (models/llama/modeling_llama.hpp:76-124), TinyLLaMAModel
//        (models/tinyllama/modeling_tinyllama.hpp:47-89), QWenForCausalLM (models/qwen/modeling_qwen.hpp:131-180),
//        weights from the same random-init .mllm file the CUDA backend loads.
// Nothing under mllm_b200/ links or loads this library.
#include <cstdint>
#include <cstring>
#include <memory>
#include <string>
#include <vector>
#include <omp.h>

#include "Backend.hpp"
#include "Generate.hpp"
#include "Layer.hpp"
#include "Module.hpp"
#include "Op.hpp"
#include "ParamLoader.hpp"
#include "Tensor.hpp"
#include "backends/cpu/CPUBackend.hpp"
#include "memory/SystemMemoryManager.hpp"
#include "models/llama/modeling_llama.hpp"
#include "models/qwen/modeling_qwen.hpp"
#include "models/tinyllama/modeling_tinyllama.hpp"

using namespace mllm;

#define API extern "C" __attribute__((visibility("default")))

static Backend *cpu_backend() {
    Module::initBackend(MLLM_CPU);
    return Backend::global_backends[MLLM_CPU].get();
}

API int cpuref_set_threads(int n) {
    if (n > 0) {
        CPUBackend::cpu_threads = n;
        omp_set_num_threads(n);
    }
    return omp_get_max_threads();
}

// ------------------------------------------------------------------ (i) ops
struct RefTensorDesc { // what Python passes for each input: memory is canonical for the ctype (BSHD: [b][s][h][d]; BCTHW: [b][c][t][h][w])
    const void *data;
    int shape[5]; // 4-D: batch, head, sequence, dimension (the reference's reshape() argument order); 5-D: batch, channel, time, height, width
    int ndim;     // 4 or 5
    int dtype;    // DataType
    int ctype;    // ChlType of the memory image (BSHD = 0, BHSD = 1, BCTHW = 3 ...)
};

struct RefOp {
    Op *op = nullptr;
    std::vector<std::shared_ptr<Tensor>> ins, outs;
};

API void *cpuref_op_create(const char **keys, const float *vals, int n, const char *name, int threads) {
    OpParam p;
    for (int i = 0; i < n; i++) p[keys[i]] = vals[i];
    Op *op = cpu_backend()->opCreate(p, name ? name : "", threads > 0 ? threads : 4);
    if (!op) return nullptr;
    op->type() = (OpType)p["type"];
    auto *r = new RefOp;
    r->op = op;
    return r;
}

API int cpuref_op_load(void *h, const char *mllm_path) {
    auto *r = (RefOp *)h;
    ParamLoader loader(mllm_path ? mllm_path : "");
    return (int)r->op->load(loader);
}

static std::shared_ptr<Tensor> make_input(const RefTensorDesc &d, int idx) {
    auto t = std::make_shared<Tensor>(cpu_backend());
    t->setName("input" + std::to_string(idx));
    t->setDtype((DataType)d.dtype);
    if (d.ndim == 5) {
        t->setCtype((ChlType)d.ctype);
        t->reshape(d.shape[0], d.shape[1], d.shape[2], d.shape[3], d.shape[4]);
    } else {
        t->setCtype((ChlType)d.ctype);
        t->reshape(d.shape[0], d.shape[1], d.shape[2], d.shape[3]);
    }
    t->alloc();
    memcpy(t->rawHostPtr(), d.data, t->cntSize());
    return t;
}

// reshape -> setUp -> execute (test/cpu/CPUTest.hpp TEST_RESHAPE / TEST_SETUP / TEST_EXCUTE); n_out output shells
API int cpuref_op_run(void *h, const RefTensorDesc *in, int n_in, int n_out) {
    auto *r = (RefOp *)h;
    r->ins.clear();
    r->outs.clear();
    for (int i = 0; i < n_in; i++) r->ins.push_back(make_input(in[i], i));
    for (int i = 0; i < n_out; i++) {
        auto t = std::make_shared<Tensor>(cpu_backend());
        t->setName("output" + std::to_string(i));
        r->outs.push_back(t);
    }
    if (r->op->reshape(r->ins, r->outs) != MLLM_NO_ERROR) return -1;
    if (r->op->setUp(r->ins, r->outs) != MLLM_NO_ERROR) return -2;
    if (r->op->execute(r->ins, r->outs) != MLLM_NO_ERROR) return -3;
    return 0;
}

// shape of output i as (batch, head, sequence, dimension) [+ 5-D: batch, channel, time, height, width], dtype, ctype
API int cpuref_op_out_info(void *h, int i, int *shape5, int *dtype, int *ctype) {
    auto *r = (RefOp *)h;
    auto &t = r->outs[i];
    *dtype = (int)t->dtype();
    *ctype = (int)t->ctype();
    if (t->ctype() == BCTHW || t->ctype() == BTHWC || t->ctype() == BWCTH) {
        shape5[0] = t->batch(), shape5[1] = t->channel(), shape5[2] = t->time(), shape5[3] = t->height(), shape5[4] = t->width();
        return 5;
    }
    shape5[0] = t->batch(), shape5[1] = t->head(), shape5[2] = t->sequence(), shape5[3] = t->dimension(), shape5[4] = 1;
    return 4;
}

// output i as fp32 in canonical [batch][sequence][head][dimension] order (or [b][c][t][h][w]), whatever its ctype/dtype
API int cpuref_op_out_read(void *h, int i, float *dst) {
    auto *r = (RefOp *)h;
    auto &t = r->outs[i];
    size_t k = 0;
    if (t->ctype() == BCTHW || t->ctype() == BTHWC || t->ctype() == BWCTH) {
        for (int b = 0; b < t->batch(); b++)
            for (int c = 0; c < t->channel(); c++)
                for (int tt = 0; tt < t->time(); tt++)
                    for (int y = 0; y < t->height(); y++)
                        for (int x = 0; x < t->width(); x++) dst[k++] = t->dataAt<float>(b, c, tt, y, x);
        return 0;
    }
    for (int b = 0; b < t->batch(); b++)
        for (int s = 0; s < t->sequence(); s++)
            for (int hh = 0; hh < t->head(); hh++)
                for (int d = 0; d < t->dimension(); d++) {
                    if (t->dtype() == MLLM_TYPE_F16) dst[k++] = MLLM_FP16_TO_FP32(t->dataAt<mllm_fp16_t>(b, hh, s, d));
                    else dst[k++] = t->dataAt<float>(b, hh, s, d);
                }
    return 0;
}

API void cpuref_op_clear_cache(void *h) { ((RefOp *)h)->op->clearCache(); }

API void cpuref_op_destroy(void *h) {
    auto *r = (RefOp *)h;
    r->ins.clear();
    r->outs.clear();
    delete r->op;
    delete r;
}

// ------------------------------------------------------------------ (ii) model graphs on MLLM_CPU
struct RefModel {
    std::unique_ptr<Module> m;
    int vocab = 0;
};

// family: "llama2-7b" (LLaMAConfig "7B"), "tinyllama-1.1b" (TinyLLaMAConfig "1.1B"), "qwen1.5-1.8b" (QWenConfig "1.8B"),
// "llama-generic": LLaMAModel's public dimension constructor (modeling_llama.hpp:92-106) with
//   dims = {vocab, hidden, heads, kv_heads, ffn, layers, max_position_embeddings}, rope_type, rope_theta (RMSNorm eps fixed 1e-6, no bias);
// "qwen-generic": QWenForCausalLM with the same dims overwritten in a QWenConfig (qkv bias, rms_eps from the argument)
API void *cpuref_model_create(const char *family, int cache_limit, int rope_type, const int *dims, float rope_theta, float rms_eps, int threads) {
    cpuref_set_threads(threads);
    cpu_backend();
    auto *r = new RefModel;
    const std::string f = family;
    if (f == "llama2-7b") {
        LLaMAConfig config(cache_limit, "7B", (RoPEType)rope_type);
        r->vocab = config.vocab_size;
        r->m.reset(new LLaMAModel(config));
    } else if (f == "tinyllama-1.1b") {
        TinyLLaMAConfig config(cache_limit, "1.1B", (RoPEType)rope_type);
        r->vocab = config.vocab_size;
        r->m.reset(new TinyLLaMAModel(config));
    } else if (f == "qwen1.5-1.8b") {
        QWenConfig config(cache_limit, "1.8B", (RoPEType)rope_type);
        r->vocab = config.vocab_size;
        r->m.reset(new QWenForCausalLM(config));
    } else if (f == "qwen-generic") { // QWenConfig is a plain struct: start from "1.8B" and overwrite the dimensions (qkv bias, config eps)
        QWenConfig config(cache_limit, "1.8B", (RoPEType)rope_type);
        config.vocab_size = dims[0], config.hidden_size = dims[1], config.num_attention_heads = dims[2], config.num_key_value_heads = dims[3];
        config.intermediate_size = dims[4], config.num_hidden_layers = dims[5], config.max_position_embeddings = dims[6];
        config.rope_theta = rope_theta, config.rms_norm_eps = rms_eps, config.tie_embedding_words = false;
        r->vocab = config.vocab_size;
        r->m.reset(new QWenForCausalLM(config));
    } else if (f == "llama-generic") {
        LLaMANameConfig names;
        names.init((RoPEType)rope_type);
        r->vocab = dims[0];
        r->m.reset(new LLaMAModel(dims[0], dims[1], dims[2], dims[3], dims[4], dims[5], (RoPEType)rope_type, rope_theta, dims[6], cache_limit,
                                  "flash_attention_2", names, names.blk_name));
    } else {
        delete r;
        return nullptr;
    }
    return r;
}

API int cpuref_model_load(void *h, const char *path) {
    ((RefModel *)h)->m->load(path);
    return 0;
}

API int cpuref_model_vocab(void *h) { return ((RefModel *)h)->vocab; }

// One forward the way the demos do it (tokens2Input, tokenizers/Tokenizer.hpp:78-101; model({input})).
// ids: [B][S] token ids as float.  logits_out: [B][S_out][vocab] fp32 with S_out returned (1 when the graph clips to the
// last position, S otherwise -- TinyLLaMAModel keeps all positions).
API int cpuref_model_forward(void *h, const float *ids, int B, int S, float *logits_out, int max_rows) {
    auto *r = (RefModel *)h;
    Tensor in(B, 1, S, 1, cpu_backend(), true);
    in.setName("input");
    Tensor::tensor_status = TENSOR_STATIC_INIT;
    in.setTtype(INPUT_TENSOR);
    for (int b = 0; b < B; b++)
        for (int s = 0; s < S; s++) in.setDataAt<float>(b, 0, s, 0, ids[b * S + s]);
    auto out = (*r->m)({in});
    Tensor &o = out[0];
    const int So = o.sequence(), V = o.dimension();
    if (o.batch() * So > max_rows) return -So;
    size_t k = 0;
    for (int b = 0; b < o.batch(); b++)
        for (int s = 0; s < So; s++)
            for (int d = 0; d < V; d++) logits_out[k++] = o.dataAt<float>(b, 0, s, d);
    return So;
}

// Module::generate as examples/demo_qwen_batch.cpp:61-68 calls it (greedy, batch of B equal-length prompts, Module.cpp:146-205).
// out_tokens: [B][max_new] filled with the generated ids, n_out[b] = how many row b produced (stops at end_token, -1 = never).
API int cpuref_model_generate(void *h, const float *ids, int B, int S, int max_new, int end_token, unsigned *out_tokens, int *n_out) {
    auto *r = (RefModel *)h;
    Tensor in(B, 1, S, 1, cpu_backend(), true);
    in.setName("input");
    Tensor::tensor_status = TENSOR_STATIC_INIT;
    in.setTtype(INPUT_TENSOR);
    for (int b = 0; b < B; b++)
        for (int s = 0; s < S; s++) in.setDataAt<float>(b, 0, s, 0, ids[b * S + s]);
    LlmTextGeneratorOpts opt{};
    opt.max_new_tokens = (size_t)max_new;
    opt.do_sample = false;
    opt.temperature = 0.3F;
    opt.top_k = 50;
    opt.top_p = 0.F;
    auto res = r->m->generate(in, opt, end_token);
    for (int b = 0; b < B; b++) {
        n_out[b] = (int)res[b].size();
        for (size_t i = 0; i < res[b].size() && (int)i < max_new; i++) out_tokens[b * max_new + i] = res[b][i];
    }
    return 0;
}

API void cpuref_model_clear(void *h) { ((RefModel *)h)->m->clear_kvcache(); }

API void cpuref_model_destroy(void *h) { delete (RefModel *)h; }

// ------------------------------------------------------------------ (iii) the reference's samplers, drawn many times
// _LlmTextGenerateTopkSamplingMethod / _LlmTextGenerateToppSamplingMethod (Generate.cpp:45-143) seed a fresh std::mt19937 from
// std::random_device on every call, so a single draw cannot be reproduced; their DISTRIBUTION can: counts[id] = how often id came out
// of `draws` calls on the same row (mode 0: top-k on logits, mode 1: top-p on probabilities).
API int cpuref_sample_hist(int mode, const float *x, int n, float p_or_k, float temperature, int draws, unsigned *counts) {
    Tensor t(1, 1, 1, n, cpu_backend(), true);
    memcpy(t.hostPtr<float>(), x, (size_t)n * 4);
    memset(counts, 0, (size_t)n * 4);
    try {
        if (mode == 0) {
            _LlmTextGenerateTopkSamplingMethod m((int)p_or_k, temperature);
            for (int i = 0; i < draws; i++) counts[m.generate(t)]++;
        } else {
            _LlmTextGenerateToppSamplingMethod m(p_or_k, temperature);
            for (int i = 0; i < draws; i++) counts[m.generate(t)]++;
        }
    } catch (const std::exception &) {
        return -1;
    }
    return 0;
}
