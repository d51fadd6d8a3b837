// run_vlm_cuda -- the reference's own VLM graphs (models/llava/modeling_llava.hpp, models/qwen2_vl/modeling_qwen2_vl.hpp, UNCHANGED) on
// the CPU backend or on the CUDA backend, chosen by a device flag as the reference's demos choose a backend, driven the way
// examples/demo_llava.cpp:45-62 and examples/demo_qwen2_vl.cpp:47-65 drive them: prefill with {ids, image[, grid]}, argmax of the last
// position, chatPostProcessing (processor/PostProcess.hpp:10-19) shrinks the ids to the new token and empties the image tensors.
// Token ids and pixels come from a seed (no vocabulary / image files are needed).
//
//   run_vlm_cuda --model llava|qwen2vl --dims ... --write FILE      one CPU pass that asks the graph for its parameter names and writes
//                                                                    a random-init .mllm (Q4_K / Q4_0 Linears, F32 everything else)
//   run_vlm_cuda --model llava|qwen2vl --dims ... --device 0|4 --file FILE --text 8 --decode 8 [--dump logits.bin]
// prints one JSON line (tokens, vision + prefill ms, decode ms/step); --dump writes the fp32 logits of every step.
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <fstream>
#include <random>
#include <string>
#include <vector>

#include <execinfo.h>
#include <signal.h>
#include <unistd.h>
#include "cmdline.h"
#include "models/llava/modeling_llava.hpp"
#include "models/qwen2_vl/modeling_qwen2_vl.hpp"
#include "processor/PostProcess.hpp"

using namespace mllm;

// ---------------------------------------------------------------------------------------------------------------- weight file
// Answers the graph's load pass with random parameters and remembers them; save() writes the reference's .mllm layout
// (ParamLoader.cpp:150-190: magic, index length, {name, size, offset, dtype} records, data).
class RandomInitLoader final : public AbstructLoader {
public:
    RandomInitLoader(DataType linear_type, unsigned seed) : linear_(linear_type), rng_(seed) {}
    DataType getDataType(string name) override {
        const bool w = name.size() > 7 && name.compare(name.size() - 7, 7, ".weight") == 0;
        const bool linear = w && (name.find("proj") != string::npos || name.find("linear") != string::npos || name.find("lm_head") != string::npos
                                  || name.find("fc") != string::npos || name.find("qkv") != string::npos || name.find("mlp.0") != string::npos
                                  || name.find("mlp.2") != string::npos || name.find("query") != string::npos || name.find("key") != string::npos
                                  || name.find("value") != string::npos || name.find("dense") != string::npos);
        const bool conv = name.find("patch_embed") != string::npos; // patch_embedding / patch_embed.proj stay F32
        if (w && name.find("embed_tokens") != string::npos) return MLLM_TYPE_Q4_0; // the quantizer's rule for the token table
        return linear && !conv ? linear_ : MLLM_TYPE_F32;
    }
    bool load(std::shared_ptr<Tensor> t) override { return load(t.get()); }
    bool load(Tensor *t) override {
        if (getenv("VLM_TRACE")) fprintf(stderr, "load %s dtype %d bytes %zu\n", t->name().c_str(), (int)t->dtype(), t->cntSize());
        const DataType dt = t->dtype();
        const size_t bytes = t->cntSize();
        std::vector<uint8_t> buf(bytes);
        std::uniform_real_distribution<float> u(-1.f, 1.f);
        const string &n = t->name();
        if (dt == MLLM_TYPE_F32) {
            float *f = (float *)buf.data();
            const bool norm_w = (n.find("norm") != string::npos || n.find("ln_") != string::npos) && n.find(".weight") != string::npos;
            const bool lin_w = getDataType(n) != MLLM_TYPE_F32 || (linear_ == MLLM_TYPE_F32 && n.find(".weight") != string::npos && !norm_w && n.find("embedding") == string::npos && n.find("patch_embed") == string::npos);
            const float a = lin_w ? 0.05f : n.find("patch_embed") != string::npos ? 0.05f : n.find(".bias") != string::npos ? 0.02f : 0.5f;
            for (size_t i = 0; i < bytes / 4; i++) f[i] = norm_w ? 1.0f + 0.1f * u(rng_) : a * u(rng_);
        } else if (dt == MLLM_TYPE_Q4_0) {
            for (size_t b = 0; b < bytes / 18; b++) {
                const mllm_fp16_t d = MLLM_FP32_TO_FP16(0.003f + 0.003f * std::fabs(u(rng_))) /* weight std ~0.02 */;
                memcpy(&buf[b * 18], &d, 2);
                for (int i = 0; i < 16; i++) buf[b * 18 + 2 + i] = (uint8_t)((1 + rng_() % 15) | ((1 + rng_() % 15) << 4)); // nibbles 1..15: (q - 8) is centred on zero
            }
        } else if (dt == MLLM_TYPE_Q4_K) { // block_q4_K: d, dmin (fp16), 12 bytes of 6-bit scales / mins, 128 bytes of nibbles
            for (size_t b = 0; b < bytes / 144; b++) {
                // weight = d * scale6 * nibble - dmin * min6 with uniform 6-bit scales / mins and nibbles: dmin = 7.5 d centres it on zero
                const float dv = 0.00005f + 0.00005f * std::fabs(u(rng_)); // weight std ~0.02 (W ~ N(0, 0.02^2))
                const mllm_fp16_t d = MLLM_FP32_TO_FP16(dv), dm = MLLM_FP32_TO_FP16(7.5f * dv);
                memcpy(&buf[b * 144], &d, 2);
                memcpy(&buf[b * 144 + 2], &dm, 2);
                for (int i = 4; i < 144; i++) buf[b * 144 + i] = (uint8_t)(rng_() & 0xff);
            }
        } else {
            fprintf(stderr, "RandomInitLoader: dtype %d of %s is not generated\n", (int)dt, n.c_str());
            exit(3);
        }
        memcpy(t->rawHostPtr(), buf.data(), bytes);
        names_.push_back(n);
        types_.push_back(dt);
        blobs_.push_back(std::move(buf));
        return true;
    }
    void save(const string &path) {
        uint64_t index_len = 0;
        for (auto &n : names_) index_len += 4 + n.size() + 8 + 8 + 4;
        std::ofstream f(path, std::ios::binary);
        const int32_t magic = 20012;
        f.write((const char *)&magic, 4);
        f.write((const char *)&index_len, 8);
        uint64_t off = 4 + 8 + index_len;
        for (size_t i = 0; i < names_.size(); i++) {
            const int32_t ln = (int32_t)names_[i].size(), dt = (int32_t)types_[i];
            const uint64_t sz = blobs_[i].size();
            f.write((const char *)&ln, 4);
            f.write(names_[i].data(), ln);
            f.write((const char *)&sz, 8);
            f.write((const char *)&off, 8);
            f.write((const char *)&dt, 4);
            off += sz;
        }
        for (auto &b : blobs_) f.write((const char *)b.data(), (std::streamsize)b.size());
    }
    size_t count() const { return names_.size(); }

private:
    DataType linear_;
    std::mt19937 rng_;
    std::vector<string> names_;
    std::vector<DataType> types_;
    std::vector<std::vector<uint8_t>> blobs_;
};

// Module::load (Module.hpp:215-225) with the loader given instead of a ParamLoader over a file
template <typename Model>
static void load_with(Model &model, std::shared_ptr<AbstructLoader> loader) {
    model.loader = std::move(loader);
    Tensor::tensor_status = TENSOR_STATIC_INIT;
    model.doLoad = true;
    model.doTrace = true;
    vector<Tensor> tmps;
    for (int i = 0; i < 5; i++) {
        Tensor t(Backend::global_backends[MLLM_CPU].get());
        t.setName("input" + std::to_string(i));
        t.reshape(1, 1, 1, 10);
        t.alloc();
        memset(t.rawHostPtr(), 0, t.cntSize());
        t.setModule(&model);
        tmps.push_back(t);
    }
    Module::llm_model_ptr = &model;
    model(tmps);
    model.doLoad = false;
    model.tracedFlag = true;
    model.doTrace = false;
}

static Tensor host_tensor(const string &name, int b, int h, int s, int d) {
    Tensor t(b, h, s, d, Backend::global_backends[MLLM_CPU].get(), true);
    t.setName(name);
    Tensor::tensor_status = TENSOR_STATIC_INIT;
    t.setTtype(INPUT_TENSOR);
    return t;
}

struct StepLog {
    int fused_steps = 0; // forwards answered by the CUDA backend's graph-replayed decode plan
    void count_fused(BackendType device) {
#ifdef USE_CUDA
        if (device == MLLM_CUDA) fused_steps += static_cast<CUDABackend *>(Backend::global_backends[MLLM_CUDA].get())->last_forward_fused;
#endif
    }
    std::vector<unsigned> tokens;
    std::vector<double> ms;
    double pool_ms0 = 0; // driver time of the memory pool before the request (start-up, weight load)
    void mark_pool(BackendType device) {
#ifdef USE_CUDA
        if (device == MLLM_CUDA) pool_ms0 = static_cast<CUDABackend *>(Backend::global_backends[MLLM_CUDA].get())->memory()->driver_ms;
#endif
    }
    FILE *dump = nullptr;
    unsigned pick(Tensor lg) { // argmax of the last position, first maximum wins (tokenizers/Tokenizer.hpp:160-170)
        if (lg.backend()->type() != MLLM_CPU) lg.cpu();
        const int s = lg.sequence() - 1, n = lg.dimension();
        std::vector<float> row(n);
        for (int i = 0; i < n; i++) row[i] = lg.dataAt<float>(0, 0, s, i);
        if (dump) fwrite(row.data(), 4, n, dump);
        return (unsigned)(std::max_element(row.begin(), row.end()) - row.begin());
    }
};

static void print_json(const char *model, int device, int prompt_tokens, const StepLog &log, const Module &m) {
    double dec = 0;
    for (size_t i = 1; i < m.inference_times_.size(); i++) dec += m.inference_times_[i];
    std::vector<double> tail(m.inference_times_.begin() + std::min<size_t>(m.inference_times_.size(), 1), m.inference_times_.end());
    std::sort(tail.begin(), tail.end());
    printf("{\"model\": \"%s\", \"device\": %d, \"prompt_tokens\": %d, \"decode\": %zu, \"prefill_ms\": %.3f, \"decode_ms_per_step\": %.4f, \"decode_ms_median\": %.4f, "
           "\"decode_ms_max\": %.3f, \"fused_steps\": %d, ", model, device, prompt_tokens, log.tokens.size(), m.inference_times_.empty() ? 0.0 : m.inference_times_[0],
           m.inference_times_.size() > 1 ? dec / (m.inference_times_.size() - 1) : 0.0, tail.empty() ? 0.0 : tail[tail.size() / 2], tail.empty() ? 0.0 : tail.back(), log.fused_steps);
#ifdef USE_CUDA
    if (device == MLLM_CUDA) { // the memory pool's driver traffic: allocations (and device-page population) made while the process ran
        auto *mm = static_cast<CUDABackend *>(Backend::global_backends[MLLM_CUDA].get())->memory();
        printf("\"pool_mb\": %.0f, \"pool_driver_calls\": %zu, \"pool_driver_ms\": %.1f, \"pool_driver_ms_in_request\": %.1f, ", mm->bytes_allocated / 1048576.0, mm->driver_calls,
               mm->driver_ms, mm->driver_ms - log.pool_ms0);
    }
#endif
    printf("\"tokens\": [");
    for (size_t i = 0; i < log.tokens.size(); i++) printf("%s%u", i ? ", " : "", log.tokens[i]);
    printf("]}\n");
}

static void on_segv(int) { void *bt[64]; int n = backtrace(bt, 64); backtrace_symbols_fd(bt, n, 2); _exit(139); }
int main(int argc, char **argv) {
    signal(SIGSEGV, on_segv);
    cmdline::parser cp;
    cp.add<string>("model", 'm', "llava | qwen2vl", false, "llava");
    cp.add<int>("device", 'd', "0 = CPU backend, 4 = CUDA backend", false, 0);
    cp.add<string>("file", 'f', ".mllm weights", false, "");
    cp.add<string>("write", 'w', "write a random-init .mllm for this model and exit", false, "");
    cp.add<string>("dims", 0, "llava: vocab,hidden,heads,ffn,layers,vhidden,vheads,vffn,vlayers,patch,img   qwen2vl: vocab,hidden,heads,kvheads,ffn,layers,vembed,grid_h,grid_w,sec0,sec1,sec2", false, "");
    cp.add<int>("text", 0, "text tokens around the image", false, 8);
    cp.add<int>("decode", 0, "greedy steps (the first comes from the prefill)", false, 8);
    cp.add<int>("limit", 'l', "KV cache size", false, 1024);
    cp.add<int>("thread", 't', "CPU backend threads", false, 4);
    cp.add<int>("seed", 's', "input seed", false, 7);
    cp.add<string>("linear", 0, "--write: dtype of the Linear weights: default (llava q4_k, qwen2vl q4_0) | f32 | q4_0 | q4_k", false, "default");
    cp.add<int>("perturb", 0, "move every pixel to the next representable float (1 ulp): the noise-floor experiment", false, 0);
    cp.add<string>("dump", 0, "write every step's last-position logits (fp32) here", false, "");
    cp.parse_check(argc, argv);
    const string which = cp.get<string>("model");
    const BackendType device = (BackendType)cp.get<int>("device");
    CPUBackend::cpu_threads = cp.get<int>("thread");
    std::vector<int> dims;
    {
        const string s = cp.get<string>("dims");
        for (size_t i = 0; i < s.size();) {
            size_t j = s.find(',', i);
            if (j == string::npos) j = s.size();
            dims.push_back(std::stoi(s.substr(i, j - i)));
            i = j + 1;
        }
    }
    auto linear_type = [&](DataType dflt) {
        const string l = cp.get<string>("linear");
        return l == "f32" ? MLLM_TYPE_F32 : l == "q4_0" ? MLLM_TYPE_Q4_0 : l == "q4_k" ? MLLM_TYPE_Q4_K : dflt;
    };
    const int n_text = cp.get<int>("text"), n_decode = cp.get<int>("decode"), limit = cp.get<int>("limit");
    std::mt19937 rng((unsigned)cp.get<int>("seed"));
    std::uniform_real_distribution<float> upix(-1.f, 1.f);
    StepLog log;
    if (!cp.get<string>("dump").empty()) log.dump = fopen(cp.get<string>("dump").c_str(), "wb");

    if (which == "llava") {
        // LLaVA-1.5-7B by default (configuration_llava.hpp:24-42: CLIP ViT-L/14-336, 23 of its 24 blocks, LLaMA-2-7B body)
        LLaVAConfig cfg(limit, "7B", 32064);
        if (dims.empty()) dims = {cfg.vocab_size, cfg.hidden_dim, cfg.head_size, cfg.ffn_hidden, cfg.block_num, cfg.vision_hidden_dim, cfg.vision_head_size, cfg.vision_ffn_hidden, cfg.vision_block_num, cfg.patch, cfg.img_hw};
        if (dims.size() != 11) { fprintf(stderr, "llava --dims needs 11 numbers\n"); return 2; }
        const int vocab = dims[0], img = dims[10], patch = dims[9];
        LLaVAModel model(vocab, dims[1], dims[2], dims[3], dims[4], cfg.RoPE_type, cfg.rope_theta, cfg.max_position_embeddings, limit, cfg.names_config, dims[5], dims[6],
                         dims[7], patch, img, dims[8], cfg.attn_implementation, cfg.vit_names_config);
        if (!cp.get<string>("write").empty()) {
            auto loader = std::make_shared<RandomInitLoader>(linear_type(MLLM_TYPE_Q4_K), 1234u);
            load_with(model, loader);
            loader->save(cp.get<string>("write"));
            printf("{\"written\": \"%s\", \"tensors\": %zu}\n", cp.get<string>("write").c_str(), loader->count());
            return 0;
        }
        if (device != MLLM_CPU) model.to(device);
        model.load(cp.get<string>("file"));
        const int img_token = 32000 < vocab ? 32000 : vocab - 1; // modeling_llava.hpp:130 looks for id 32000
        const int S = n_text + 1, img_pos = n_text / 2;
        Tensor ids = host_tensor("input_ids", 1, 1, S, 1);
        for (int s = 0; s < S; s++) {
            unsigned v = rng() % (unsigned)vocab;
            if ((int)v == img_token) v = 1;
            ids.setDataAt<float>(0, 0, s, 0, s == img_pos ? (float)img_token : (float)v);
        }
        if (img_token != 32000) { fprintf(stderr, "llava: the graph looks for token 32000; vocab must exceed it\n"); return 2; }
        Tensor image = host_tensor("input_imgs", 1, img, 3, img); // [B, head = H, seq = C, dim = W]
        for (int c = 0; c < 3; c++)
            for (int h = 0; h < img; h++)
                for (int w = 0; w < img; w++) { const float v = upix(rng); image.setDataAt<float>(0, h, c, w, cp.get<int>("perturb") ? std::nextafterf(v, 2.f) : v); }
        const int prompt_tokens = S - 1 + (img / patch) * (img / patch);
        log.mark_pool(device);
        for (int step = 0; step < n_decode; step++) {
            auto result = model({ids, image});
            log.count_fused(device);
            const unsigned tok = log.pick(result[0]);
            log.tokens.push_back(tok);
            chatPostProcessing(tok, ids, {&image});
        }
        print_json("llava", (int)device, prompt_tokens, log, model);
    } else if (which == "qwen2vl") {
        // Qwen2-VL-2B by default (configuration_qwen.hpp "1.5b-vl" + configuration_qwen2_vl.hpp:40-60; 32 vision blocks of 16 heads)
        Qwen2VLConfig cfg(limit, "1.5b-vl");
        int gh = 74, gw = 74; // 1036 x 1036 pixels = 74 x 74 patches of 14 -> 1369 image tokens (BASELINE configs[3])
        if (!dims.empty()) {
            if (dims.size() != 12) { fprintf(stderr, "qwen2vl --dims needs 12 numbers\n"); return 2; }
            cfg.vocab_size = dims[0]; cfg.hidden_size = dims[1]; cfg.num_attention_heads = dims[2]; cfg.num_key_value_heads = dims[3];
            cfg.intermediate_size = dims[4]; cfg.num_hidden_layers = dims[5]; cfg.vision_embed_dim = dims[6]; gh = dims[7]; gw = dims[8];
            cfg.mrope_section = {dims[9], dims[10], dims[11]};
        }
        cfg.cache_limit = limit;
        Qwen2VLModel model(cfg);
        if (!cp.get<string>("write").empty()) {
            auto loader = std::make_shared<RandomInitLoader>(linear_type(MLLM_TYPE_Q4_0), 4321u);
            load_with(model, loader);
            loader->save(cp.get<string>("write"));
            printf("{\"written\": \"%s\", \"tensors\": %zu}\n", cp.get<string>("write").c_str(), loader->count());
            return 0;
        }
        if (device != MLLM_CPU) model.to(device);
        model.load(cp.get<string>("file"));
        const int vocab = cfg.vocab_size;
        if (cfg.image_token_id >= vocab || cfg.vision_start_token_id >= vocab) { fprintf(stderr, "qwen2vl: vocab must exceed the vision token ids\n"); return 2; }
        const int n_patch = gh * gw, n_img = n_patch / (cfg.spatial_merge_size * cfg.spatial_merge_size);
        const int pre = n_text / 2, S = n_text + 2 + n_img;
        Tensor ids = host_tensor("input_ids", 1, 1, S, 1);
        for (int s = 0; s < S; s++) {
            unsigned v = rng() % 1000u; // plain text ids, far from the special tokens
            float id = (float)v;
            if (s == pre) id = (float)cfg.vision_start_token_id;
            else if (s > pre && s <= pre + n_img) id = (float)cfg.image_token_id;
            else if (s == pre + n_img + 1) id = (float)cfg.vision_end_token_id;
            ids.setDataAt<float>(0, 0, s, 0, id);
        }
        // pixel patches as the processor hands them over: BCTHW [patches, 3, 2, 14, 14] (processing_qwen2_vl.hpp), grid (t, h, w)
        Tensor patches(Backend::global_backends[MLLM_CPU].get());
        patches.setName("pixel_values");
        patches.reshape(n_patch, 3, 2, 14, 14);
        patches.setDtype(MLLM_TYPE_F32);
        patches.alloc();
        patches.setTtype(INPUT_TENSOR);
        for (size_t i = 0; i < (size_t)n_patch * 3 * 2 * 14 * 14; i++) { const float v = upix(rng); patches.hostPtr<float>()[i] = cp.get<int>("perturb") ? std::nextafterf(v, 2.f) : v; }
        Tensor grid = host_tensor("image_grid_thw", 1, 1, 1, 3);
        grid.setDataAt<float>(0, 0, 0, 0, 1.f);
        grid.setDataAt<float>(0, 0, 0, 1, (float)gh);
        grid.setDataAt<float>(0, 0, 0, 2, (float)gw);
        vector<Tensor> inputs = {ids, patches, grid};
        log.mark_pool(device);
        for (int step = 0; step < n_decode; step++) {
            model.get_position_ids(inputs);
            auto result = model(inputs);
            log.count_fused(device);
            const unsigned tok = log.pick(result[0]);
            log.tokens.push_back(tok);
            chatPostProcessing(tok, inputs[0], {&inputs[1], &inputs[2]});
        }
        print_json("qwen2vl", (int)device, S, log, model);
    } else {
        fprintf(stderr, "unknown --model %s\n", which.c_str());
        return 2;
    }
    if (log.dump) fclose(log.dump);
    return 0;
}
