// run_llama_cuda -- the reference's own model graphs (models/llama/modeling_llama.hpp, models/tinyllama/modeling_tinyllama.hpp,
// models/qwen/modeling_qwen.hpp, UNCHANGED) on the CPU backend or on the CUDA backend, chosen by a device flag exactly as the
// reference's demos choose OpenCL (examples/demo_qwen.cpp:42-58: `model = model.to((BackendType)device); model.load(path)`).
// Token ids come from a seed instead of a tokenizer (no vocabulary files are needed), the greedy loop is the demos':
// model({input}) -> argmax of the last position -> feed back (examples/demo_llama.cpp:40-50, tokenizers/Tokenizer.hpp:160-170).
//
//   run_llama_cuda --device 0|4 --model llama2-7b|tinyllama-1.1b|qwen1.5-1.8b --file m.mllm --prompt 128 --decode 32 [--limit 512]
// prints one JSON line: tokens, prefill / decode tok/s from Module::profiling's inference_times_.
#include <algorithm>
#include <chrono>
#include <cstdio>
#include <cstring>
#include <random>
#include <string>
#include <vector>

#include "cmdline.h"
#include "models/llama/modeling_llama.hpp"
#include "models/qwen/modeling_qwen.hpp"
#include "models/tinyllama/modeling_tinyllama.hpp"

using namespace mllm;

static Tensor make_input(const std::vector<unsigned> &ids, int B, int S) {
    Tensor t(B, 1, S, 1, Backend::global_backends[MLLM_CPU].get(), true);
    t.setName("input");
    Tensor::tensor_status = TENSOR_STATIC_INIT;
    t.setTtype(INPUT_TENSOR);
    for (int b = 0; b < B; b++)
        for (int s = 0; s < S; s++) t.setDataAt<float>(b, 0, s, 0, (float)ids[(size_t)b * S + s]);
    return t;
}

template <typename Model>
static int run(Model &model, BackendType device, const std::string &file, int vocab, int B, int S, int n_decode, unsigned seed, int repeat = 1) {
    if (device != MLLM_CPU) model.to(device);
    model.load(file);
    std::mt19937 rng(seed);
    std::vector<unsigned> ids((size_t)B * S);
    for (auto &v : ids) v = rng() % vocab;
    std::vector<std::vector<unsigned>> out(B);
    int fused_steps = 0;
    std::vector<double> pre_all, med_all; // --repeat: the same request again after clear_kvcache() (first-request costs -- lazy kernel
                                          // loading, host-built RoPE tables, pool growth -- show up in entry 0 only)
    double pre_ms = 0.0, dec_ms = 0.0, dec_median = 0.0;
    for (int rep = 0; rep < repeat; rep++) {
    if (rep) { model.clear_kvcache(); model.inference_times_.clear(); }
    Tensor in = make_input(ids, B, S);
    for (int step = 0; step < n_decode; step++) {
        auto result = model({in});
#ifdef USE_CUDA
        if (device == MLLM_CUDA && rep == 0) fused_steps += static_cast<CUDABackend *>(Backend::global_backends[MLLM_CUDA].get())->last_forward_fused;
#endif
        Tensor lg = result[0];
        if (lg.backend()->type() != MLLM_CPU) lg.cpu(); // as the OpenCL demos do before reading a result on the host
        std::vector<unsigned> next(B);
        for (int b = 0; b < B; b++) {
            const int s = lg.sequence() - 1;
            int best = 0;
            float bv = lg.dataAt<float>(b, 0, s, 0);
            for (int i = 1; i < lg.dimension(); i++) {
                const float v = lg.dataAt<float>(b, 0, s, i);
                if (v > bv) { bv = v; best = i; }
            }
            next[b] = (unsigned)best;
            if (rep == 0) out[b].push_back((unsigned)best);
        }
        in = make_input(next, B, 1);
    }
    const double p_ms = model.inference_times_.empty() ? 0.0 : model.inference_times_[0];
    double d_ms = 0.0;
    for (size_t i = 1; i < model.inference_times_.size(); i++) d_ms += model.inference_times_[i];
    std::vector<double> tail(model.inference_times_.begin() + std::min<size_t>(model.inference_times_.size(), rep ? 1 : 4), model.inference_times_.end());
    std::sort(tail.begin(), tail.end());
    const double med = tail.empty() ? 0.0 : tail[tail.size() / 2];
    pre_all.push_back(p_ms);
    med_all.push_back(med);
    if (rep == 0) { pre_ms = p_ms; dec_ms = d_ms; dec_median = med; }
    }
    printf("{\"device\": %d, \"B\": %d, \"prompt\": %d, \"decode\": %d, \"prefill_ms\": %.3f, \"decode_ms_per_step\": %.4f, \"decode_ms_median\": %.4f, \"fused_steps\": %d, ", (int)device, B,
           S, n_decode, pre_ms, n_decode > 1 ? dec_ms / (n_decode - 1) : 0.0, dec_median, fused_steps);
    if (repeat > 1) {
        printf("\"prefill_ms_per_request\": [");
        for (size_t i = 0; i < pre_all.size(); i++) printf("%s%.3f", i ? ", " : "", pre_all[i]);
        printf("], \"decode_ms_median_per_request\": [");
        for (size_t i = 0; i < med_all.size(); i++) printf("%s%.4f", i ? ", " : "", med_all[i]);
        printf("], ");
    }
    printf("\"tokens\": [");
    for (int b = 0; b < B; b++) {
        printf("%s[", b ? ", " : "");
        for (size_t i = 0; i < out[b].size(); i++) printf("%s%u", i ? ", " : "", out[b][i]);
        printf("]");
    }
    printf("]}\n");
    return 0;
}

int main(int argc, char **argv) {
    cmdline::parser cmd;
    cmd.add<int>("device", 'd', "0 = MLLM_CPU, 4 = MLLM_CUDA", false, 0);
    cmd.add<std::string>("model", 'n', "llama2-7b | tinyllama-1.1b | qwen1.5-1.8b", false, "tinyllama-1.1b");
    cmd.add<std::string>("file", 'm', ".mllm weight file", true, "");
    cmd.add<int>("prompt", 'i', "prompt tokens", false, 32);
    cmd.add<int>("decode", 'p', "greedy steps", false, 16);
    cmd.add<int>("batch", 'b', "batch", false, 1);
    cmd.add<int>("limit", 'l', "KV cache limit", false, 512);
    cmd.add<int>("thread", 't', "CPU threads", false, 4);
    cmd.add<int>("seed", 's', "token seed", false, 7);
    cmd.add<int>("repeat", 'r', "run the request this many times (clear_kvcache between); times per request are reported", false, 1);
    cmd.add<std::string>("dims", 'x', "llama-generic: vocab,hidden,heads,kv_heads,ffn,layers,rope_type(2|4)", false, "");
    cmd.parse_check(argc, argv);
    const auto device = (BackendType)cmd.get<int>("device");
    assert(device == MLLM_CPU || device == MLLM_CUDA);
    CPUBackend::cpu_threads = cmd.get<int>("thread");
    const std::string name = cmd.get<std::string>("model"), file = cmd.get<std::string>("file");
    const int S = cmd.get<int>("prompt"), n = cmd.get<int>("decode"), B = cmd.get<int>("batch"), limit = cmd.get<int>("limit");
    const unsigned seed = (unsigned)cmd.get<int>("seed");
    const int repeat = std::max(1, cmd.get<int>("repeat"));
    Module::initBackend(MLLM_CPU);
    if (name == "llama2-7b") {
        LLaMAConfig config(limit, "7B", LLAMAROPE);
        LLaMAModel model(config);
        return run(model, device, file, config.vocab_size, B, S, n, seed, repeat);
    }
    if (name == "tinyllama-1.1b") {
        TinyLLaMAConfig config(limit, "1.1B", HFHUBROPE);
        TinyLLaMAModel model(config);
        return run(model, device, file, config.vocab_size, B, S, n, seed, repeat);
    }
    if (name == "qwen1.5-1.8b") {
        QWenConfig config(limit, "1.8B", RoPEType::HFHUBROPE);
        QWenForCausalLM model(config);
        return run(model, device, file, config.vocab_size, B, S, n, seed, repeat);
    }
    if (name == "llama-generic") { // LLaMAModel's public dimension constructor (modeling_llama.hpp:92-106): small graphs for exact checks
        int d[7] = {512, 256, 4, 4, 512, 2, 2};
        sscanf(cmd.get<std::string>("dims").c_str(), "%d,%d,%d,%d,%d,%d,%d", &d[0], &d[1], &d[2], &d[3], &d[4], &d[5], &d[6]);
        LLaMANameConfig names;
        names.init((RoPEType)d[6]);
        LLaMAModel model(d[0], d[1], d[2], d[3], d[4], d[5], (RoPEType)d[6], 10000.f, limit, limit, "flash_attention_2", names, names.blk_name);
        return run(model, device, file, d[0], B, S, n, seed, repeat);
    }
    fprintf(stderr, "unknown model %s\n", name.c_str());
    return 2;
}
