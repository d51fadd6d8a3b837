/* model_runner.h -- CPU LLaMA-family forward used as the model-level oracle and
 * as the timed CPU baseline.  TEST INFRASTRUCTURE ONLY (see oracle.c header).
 *
 * The graph is the one mllm's LLaMAModel / MultiHeadAttention / LLaMAMLP build
 * (models/llama/modeling_llama.hpp:29-115, models/transformer/
 * modeling_transformer.hpp:132-212): embedding -> N x [RMSNorm, q/k/v Linear,
 * RoPE(q) fp32, RoPE(k) -> fp16 cache, v -> fp16 cache, flash-attention-2,
 * o Linear, +residual, RMSNorm, gate/up Linear, SiLU, mul, down Linear,
 * +residual] -> RMSNorm -> last token -> lm_head.
 *
 * Compiled twice: -DRUNNER_PORT (kernels = oracle.c's orc_*) into liboracle.so
 * and -DRUNNER_REF (Linear / SiLU / attention = the reference's own compiled
 * ggml + FlashAttention2 bodies via ref_shim.cpp) into _ref/libmllm_ref.so.
 * RMSNorm, RoPE and the embedding gather are oracle.c's restatements in both,
 * because the reference has them only as Op classes bound to its Tensor
 * machinery (op/CPURMSNorm.cpp, op/CPURoPE.cpp, op/CPUEmbedding.cpp). */
#ifndef ORACLE_MODEL_RUNNER_H
#define ORACLE_MODEL_RUNNER_H
#include <stddef.h>
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

enum { MR_ATTN_NORM = 0, MR_WQ, MR_WK, MR_WV, MR_WO, MR_BQ, MR_BK, MR_BV, MR_FFN_NORM, MR_W_GATE, MR_W_UP, MR_W_DOWN, MR_N_ROLES };
enum { MR_G_EMBED = 0, MR_G_NORM, MR_G_LM_HEAD, MR_N_GLOBAL };

/* cfg_i: vocab, hidden, layers, heads, kv_heads, head_dim, ffn, rope_type(2=llama,4=hf),
 *        cache_limit, max_batch, wtype(2=Q4_0), pos_max ; cfg_f: rope_theta, rms_eps */
void *mr_create(const int *cfg_i, const float *cfg_f);
void mr_destroy(void *h);
void mr_set_layer(void *h, int layer, int role, const void *ptr);
void mr_set_global(void *h, int role, const void *ptr);
void mr_reset(void *h);
int mr_cache_len(void *h);
/* ids: [B,S] fp32 token ids.  logits: [B,vocab] for the last position.
 * dump (optional): [layers+1][B*S][hidden] hidden state after embedding and each block. */
int mr_forward(void *h, const float *ids, int B, int S, float *logits, float *dump);
#ifdef __cplusplus
}
#endif
#endif
