/*
 * mllmcu.h -- the thin C ABI under mllm's CUDA backend (libmllmcu.so).
 *
 * This is the drop-in boundary for the quantized-inference hot path: the
 * entry points below are exactly what `mllm/backends/cuda/CUDA*.cpp` op classes
 * call from their Op::execute() (and what the tests bind with ctypes).  Each
 * one names the reference CPU function it replaces (paths relative to
 * /root/reference/mllm).  Conventions:
 *   - every pointer is a DEVICE pointer owned by the caller unless it says host;
 *   - no hidden allocation, no global state: scratch memory is passed in
 *     explicitly (`ws`, sized by the matching *_workspace_bytes());
 *   - `stream` is a cudaStream_t passed as void*; all work is asynchronous on it;
 *   - return value: 0 = ok, otherwise a cudaError_t value (>0) or a negative
 *     MLLMCU_E* argument error.  Nothing falls back to the CPU.
 * Activations are fp32 in mllm's BSHD layout (Types.hpp:99-112); weights are
 * the `.mllm` file's native blocks, uploaded byte-for-byte (DataType.hpp:75-163).
 */
#ifndef MLLMCU_H
#define MLLMCU_H
#include <stddef.h>
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

#define MLLMCU_EINVAL (-1)   /* bad shape / alignment */
#define MLLMCU_ENOSUP (-2)   /* dtype or size not supported by this build */
#define MLLMCU_EWS (-3)      /* workspace too small */

/* DataType values, identical to mllm/Types.hpp:64-74 */
#define MLLMCU_TYPE_F32 0
#define MLLMCU_TYPE_F16 1
#define MLLMCU_TYPE_Q4_0 2
#define MLLMCU_TYPE_Q8_0 8
#define MLLMCU_TYPE_Q4_K 12
#define MLLMCU_TYPE_Q6_K 14
#define MLLMCU_TYPE_Q4_0_4_4 19
/* RoPEType values, identical to mllm/Types.hpp RoPEType */
#define MLLMCU_ROPE_LLAMA 2
#define MLLMCU_ROPE_HF 4

int mllmcu_version(void);
int mllmcu_device_count(void);
int mllmcu_set_device(int dev);
/* name: >=256 bytes host buffer; sm_count/mem_bytes host out-params */
int mllmcu_device_info(int dev, char *name, int *sm_major, int *sm_minor, int *sm_count, size_t *mem_bytes);

/* ---- memory: what Backend::alloc_device / free_device / copy_from_host /
 * copy_to_host (Backend.hpp:60-71) and load_from_file (:118) reduce to. */
int mllmcu_malloc(void **dptr, size_t bytes);
int mllmcu_free(void *dptr);
/* Unified (managed) memory made resident on `device` at creation (preferred location = device, prefetched): what the in-tree
 * backend's MemoryManager hands to TensorImpl::host_ptr_, so the reference's pointer-arithmetic tensor views work on device
 * memory and host code may still dereference a tensor (integration/mllm/backends/cuda/CUDABackend.hpp).  Freed with mllmcu_free. */
int mllmcu_malloc_managed(void **ptr, size_t bytes, int device);
/* Bring a range of unified memory (back) to `device` after host code wrote it -- CPUKVCache zero-fills its cache with memset at
 * construction (op/CPUKVCache.cpp:55-66, :85-96), which moves those pages to the host; without this every first kernel write to a cache
 * page faults.  A hint: ranges that are not managed memory are left alone. */
int mllmcu_mem_prefetch(void *ptr, size_t bytes, int device, void *stream);
int mllmcu_memset(void *dptr, int value, size_t bytes, void *stream);
int mllmcu_memcpy_h2d(void *dst, const void *host_src, size_t bytes, void *stream);
int mllmcu_memcpy_d2h(void *host_dst, const void *src, size_t bytes, void *stream);
int mllmcu_memcpy_d2d(void *dst, const void *src, size_t bytes, void *stream);
int mllmcu_host_alloc(void **hptr, size_t bytes); /* pinned host memory */
int mllmcu_host_free(void *hptr);
int mllmcu_stream_create(void **stream);
int mllmcu_stream_destroy(void *stream);
int mllmcu_stream_sync(void *stream);
/* CUDA events recorded on the caller's stream: device-side timing (what Module::profiling's wall
 * clock, mllm/Module.cpp:25-61, becomes on an asynchronous backend) */
int mllmcu_event_create(void **ev);
int mllmcu_event_destroy(void *ev);
int mllmcu_event_record(void *ev, void *stream);
int mllmcu_event_sync(void *ev);
int mllmcu_event_elapsed_ms(void *start, void *stop, float *ms);

/* ---- CUDA graphs.  A decode step is ~200 small launches; the backend captures the step once
 * (begin_capture .. end_capture on its stream, thread-local mode) and replays it per token.  The
 * reference has no counterpart: its CPU ops are called directly (CPUBackend::runOp, CPUBackend.cpp:314-405). */
int mllmcu_graph_begin_capture(void *stream);
int mllmcu_graph_end_capture(void *stream, void **graph_exec);
int mllmcu_graph_abort_capture(void *stream);
int mllmcu_graph_launch(void *graph_exec, void *stream);
int mllmcu_graph_destroy(void *graph_exec);

/* ---- quantize_row_q8_0 (ggml/QuantizeQ8.cpp:57-189, AVX2 body) for `rows`
 * fp32 rows of length k (row stride ldx floats) -> rows x k/32 native
 * block_q8_0 (34 B each).  Bytes are identical to the reference's. */
int mllmcu_quantize_q8_0(const float *x, void *y_blocks, int rows, int k, int ldx, void *stream);

/* ---- the integer part of vec_dot_q4_0_q8_0 (ggml/VecDotQ4.cpp:523-543):
 * out[n][m][kb] = sum_{e<32} (w[n][kb].q[e]-8) * x[m][kb].q[e]  (int32, exact).
 * W: N rows of K/32 block_q4_0; xq: M rows of K/32 block_q8_0.  Test/diagnostic
 * entry point proving the block dot products are bit-exact. */
int mllmcu_block_isums_q4_0_q8_0(const void *W, const void *xq_blocks, int32_t *out, int M, int N, int K, void *stream);

/* ---- Linear: CPULinear::execute -> mat_mul (op/CPULinear.cpp:97-233,
 * compute/Matmul.cpp:19-279).  y[m][n] = sum_k x[m][k] * W[n][k] + bias[n],
 * computed as the reference does: x rows quantised to Q8_0 (bit-identical),
 * integer block dots, fp32 scale-accumulate; bias added in fp32; result stored
 * fp32 or (out_f16) fp16 RNE -- the latter is how v_proj writes into the KV
 * cache slice (Matmul.cpp:262-268).
 *   wtype: MLLMCU_TYPE_Q4_0 | MLLMCU_TYPE_Q8_0;  W: [N][K/32] native blocks,
 *   x: [M][K] fp32 row stride ldx;  y: [M][N] row stride ldy (elements).
 * wtype MLLMCU_TYPE_Q4_K | MLLMCU_TYPE_Q6_K (K % 256 == 0): the k-quant path -- rows quantised to block_q8_K
 * (quantize_row_q8_K, ggml/QuantizeQ8.cpp:216-266, bytes identical), vec_dot_q4_K_q8_K / vec_dot_q6_K_q8_K
 * (ggml/VecDotQ4.cpp:32-345, ggml/VecDotQ6.cpp:32-262): per super-block the same two exact integers
 * (sum scale*dot, sum min*bsum) times d = d_x*fp16(d_w), dmin = -d_x*fp16(dmin_w).  W: [N][K/256] block_q4_K (144 B,
 * 16-byte aligned rows) or block_q6_K (210 B).
 * ws: scratch of at least mllmcu_linear_workspace_bytes(M,K). */
size_t mllmcu_linear_workspace_bytes(int M, int K);
int mllmcu_linear(int wtype, const void *W, const float *x, const float *bias, void *y, int M, int N, int K, int ldx,
                  int ldy, int out_f16, void *ws, size_t ws_bytes, void *stream);
/* The same Linear when several Linears read ONE activation tensor back to back (q/k/v, gate/up: models/llama/
 * modeling_llama.hpp:29-36, models/transformer/modeling_transformer.hpp:150-160): the reference's mat_mul re-quantises the
 * row for every call (compute/Matmul.cpp:104-115); here flags bit 0 says "ws still holds the quantised image of exactly
 * these activations (same x, M, K, ldx) written by the previous call", so the quantiser launch is skipped.  Only valid when
 * mllmcu_linear_quant_kind() returned the same non-zero kind for both calls (0: this shape keeps no reusable image). */
int mllmcu_linear_quant_kind(int wtype, int M, int N, int K, const void *W);
int mllmcu_linear_ex(int wtype, const void *W, const float *x, const float *bias, void *y, int M, int N, int K,
                     int ldx, int ldy, int out_f16, void *ws, size_t ws_bytes, int flags, void *stream);
/* mllmcu_linear without the workspace argument, for callers that cannot call cudaMalloc (code under CUDA graph capture,
 * tight serving loops, harnesses that forbid allocation).  No counterpart in the reference.
 * The scratch is a file-scope static __device__ array of MLLMCU_LINEAR_STATIC_WS_BYTES inside the
 * library (every GPU that loads the library holds it until exit); the call allocates nothing.  Shapes whose
 * mllmcu_linear_workspace_bytes(M, K) exceeds the array return MLLMCU_EWS (and say so on stderr).  All calls on a device
 * share the array: calls that overlap on the GPU (two streams) race on it; back-to-back calls on one stream are safe.
 * Results are those of mllmcu_linear (same kernels). */
#define MLLMCU_LINEAR_STATIC_WS_BYTES (16u << 20)
size_t mllmcu_linear_static_ws_capacity(void);
int mllmcu_linear_static_ws(int wtype, const void *W, const float *x, const float *bias, void *y, int M, int N, int K,
                            int ldx, int ldy, int out_f16, void *stream);

/* ---- Fused decode Linear (M <= 8 rows): one launch = [RMSNorm ->] Q8_0 quantise -> nW (<=3) block-quantised
 * Linears sharing the input -> epilogue.  It is the same arithmetic, in the same order, as the separate
 * calls mllmcu_rmsnorm -> mllmcu_linear (-> mllmcu_add | mllmcu_silu + mllmcu_mul), so results are
 * bit-identical to that sequence; the fusion only removes launches and HBM round trips of the graph
 * RMSNorm -> {q,k,v}_proj / {gate,up}_proj -> SiLU*mul, o_proj / down_proj -> +residual
 * (models/llama/modeling_llama.hpp:29-68, models/transformer/modeling_transformer.hpp:132-212).
 *   W[i]: [Ns[i]][K/32] block_q4_0 (host array of nW device pointers); bias: host array of nW device
 *   pointers (entries or the array itself may be NULL); x: [M][K] fp32, row stride ldx;
 *   norm_w: RMSNorm weight [K] or NULL (no norm); eps: RMSNorm epsilon;
 *   epilogue 0: y[m][n] = W.x (+bias), outputs of the nW matrices concatenated along n;
 *            1: same + residual[m][n] (row stride ldr);
 *            2: nW == 2, y[m][n] = silu(W0.x)[n] * (W1.x)[n];
 *   y row stride ldy, fp32 or (out_f16) fp16; pdl != 0 launches with programmatic dependent launch so the
 *   weight prefetch of this kernel overlaps the tail of the previous kernel on the stream.
 * Returns MLLMCU_ENOSUP when the shape is outside this path (caller then uses mllmcu_linear). */
int mllmcu_gemv_fused(int wtype, const void *const *W, const int *Ns, const float *const *bias, int nW, const float *x,
                      int ldx, int M, int K, const float *norm_w, float eps, int epilogue, const float *residual, int ldr,
                      void *y, int ldy, int out_f16, int pdl, void *stream);

/* mllmcu_gemv_fused described by a struct, with two extras for the chained decode step (neither changes results):
 *  - replicas: the 148 CTAs of a launch all read the same activation row, which sits in a handful of L2 slices; when the
 *    producer stores `y_reps` copies (element stride y_rep_stride) and the consumer is told `x_reps` / x_rep_stride, CTA c
 *    reads copy c % x_reps and the L2 broadcast spreads over x_reps times more slices (same for the RMSNorm weight);
 *  - pdl == 2: like pdl == 1, but the kernel lets its dependents start only after its OWN dependency wait has passed, so
 *    everything launched behind it may assume that whatever ran before it (the embedding kernel bumping the position
 *    counter) is complete even ahead of their own waits;
 *  - L2 look-ahead: while the kernel streams its own weights it asks L2 to fetch n_pf (<= 8) byte ranges (16-byte aligned,
 *    sizes multiples of 16) that LATER launches will stream, paced with its own row groups. */
typedef struct mllmcu_gemv_desc {
    int wtype, nW;
    const void *W[3];
    int N[3];
    const float *bias[3];
    const float *x;
    int ldx, M, K;
    const float *norm_w;
    float eps;
    int epilogue;
    const float *residual;
    int ldr;
    void *y;
    int ldy, out_f16, pdl;
    int x_reps, norm_reps, y_reps;             /* 0 or 1: a single copy */
    long long x_rep_stride, norm_rep_stride, y_rep_stride; /* elements between copies */
    int n_pf;
    const void *pf_ptr[8];
    size_t pf_bytes[8];
    /* tensor-parallel row-parallel Linear (o_proj / down_proj of one rank): xchg_world > 1 makes the launch
     * "partial GEMV -> all-reduce over peer memory -> + residual" in ONE kernel: the partial goes to this rank's shared region
     * (y is ignored; epilogue 0, no norm, M <= 2), and the last CTA to finish announces it, waits for the peers and writes
     * xchg_out[M][N] = xchg_residual + sum over ranks (mllmcu_p2p_allreduce's arithmetic and arguments; xchg_ticket: a device
     * word, zero before the first launch, that the kernel leaves zero). */
    int xchg_world, xchg_rank;
    const void *xchg_peer_bases[8];
    size_t xchg_buf_offset, xchg_flag_offset;
    const unsigned *xchg_epoch;
    unsigned *xchg_ticket;
    const float *xchg_residual;
    float *xchg_out;
} mllmcu_gemv_desc;
int mllmcu_gemv_fused_ex(const mllmcu_gemv_desc *d, void *stream);
/* Largest M mllmcu_gemv_fused / _ex accepts: 64 (3..64 rows run on the tensor cores, gemv_mma.cu: one mma.sync.m16n8k32.u8.s8 per
 * 16 weight rows x 8 batch rows x Q block, exact integer block sums -- the demo_qwen_batch shape, examples/demo_qwen_batch.cpp:61-68),
 * or 8 when MLLMCU_GEMV_MMA=0 keeps the CUDA-core kernel. */
int mllmcu_gemv_fused_max_rows(void);

/* development aid (MLLMCU_GEMV_PROF=1): 64 %globaltimer slots {entry, dependency wait passed, prologue done, exit, 4 spare, then per warp w: [8+2w] activation
 * loads landed, [9+2w] quantisation done} of CTA 0 for the
 * last n mllmcu_gemv_fused launches, oldest first; returns the number of launches copied */
int mllmcu_gemv_prof_read(unsigned long long *host_out, int n_launches);

/* development aid (MLLMCU_ATTN_PROF=1): 8 %globaltimer slots per mllmcu_decode_attn_rope launch, oldest first: CTA 0 {entry,
 * dependency wait passed, q rotated, key loop done, partial stored, ticket taken}, merging CTA of row 0 {merge start, output stored} */
int mllmcu_decode_attn_prof_read(unsigned long long *host_out, int n_launches);

/* Self-test of the branch-free IEEE division used for Q8_0 scales inside the fused kernels: counts fp32 bit
 * patterns amax in [lo, hi] for which (127/amax, amax/127) differ from IEEE division.  *mismatch_dev (device,
 * zero-initialised by the caller) receives the count; 0 is the only acceptable answer. */
int mllmcu_selftest_div(uint32_t lo, uint32_t hi, unsigned long long *mismatch_dev, void *stream);

/* ---- RMSNorm: CPURMSNorm::execute (op/CPURMSNorm.cpp:31-95): double sum of
 * squares, y = (x * rsqrt) * w  (or * (1+w) when add_unit_offset). */
int mllmcu_rmsnorm(const float *x, const float *w, float *y, int rows, int dim, float eps, int add_unit_offset, void *stream);
/* ---- LayerNorm: CPULayerNorm::execute (op/CPULayerNorm.cpp:49-89); w/b may be NULL */
int mllmcu_layernorm(const float *x, const float *w, const float *b, float *y, int rows, int dim, float eps, void *stream);

/* ---- RoPE: CPURoPE::rope_llama / rope_hf (op/CPURoPE.cpp:200-372) with the
 * host-built sin/cos tables ([pos_max][tdim], built as :75-128 does).  x is
 * [B][S][H][D] fp32; output element (b,s,h,d) goes to
 *   out + b*out_bstride + s*out_sstride + h*D + d   (element units)
 * as fp32 or fp16 (RNE) -- so k_rope can write straight into the cache slice
 * exactly as the reference's child-tensor aliasing does (Tensor.cpp:225-275).
 * Row s uses table row pos0 + s (h_cnt_, CPURoPE.cpp:449-515). */
int mllmcu_rope(int rope_type, const float *x, void *out, int B, int S, int H, int D, int rot_dim, int pos0,
                const float *sin_t, const float *cos_t, int tdim, int out_f16, long long out_bstride,
                long long out_sstride, void *stream);

/* ---- SiLU: mllm_vec_silu_f32 (compute/ActivationFunction.cpp:5-27, polynomial
 * exp of ActivationFunction.hpp:96-141 restated lane-exactly); and the fused
 * SiLU(gate)*up of LLaMAMLP (models/llama/modeling_llama.hpp:29-36). */
int mllmcu_silu(const float *x, float *y, size_t n, void *stream);
int mllmcu_silu_mul(const float *gate, const float *up, float *y, size_t n, void *stream);
/* ---- element-wise tensor-tensor ops: CPUaddTwoFunction / CPUmulTwoFunction (op/CPUBinaryFunc.hpp) */
int mllmcu_add(const float *a, const float *b, float *y, size_t n, void *stream);
int mllmcu_mul(const float *a, const float *b, float *y, size_t n, void *stream);

/* ---- Embedding: CPUEmbedding::execute, Q4_0 table (op/CPUEmbedding.cpp:66-81):
 * ids are fp32 (device); negative id leaves the row untouched. */
int mllmcu_embedding_q4_0(const void *table, const float *ids, float *out, int n_ids, int hidden, void *stream);
/* the same gather from an F32 table (op/CPUEmbedding.cpp:44-63; e.g. CLIP's position_embedding, models/llava/modeling_llava.hpp:51) */
int mllmcu_embedding_f32(const void *table, const float *ids, float *out, int n_ids, int hidden, void *stream);

/* ---- Flash attention 2 with fp32 Q and fp16 K/V cache:
 * CPUFlashAttention2Func::execute -> flash_attention_2_forward
 * (op/CPUFlashAttention2Func.hpp:49-125, compute/FlashAttention2.hpp:1212-2284).
 * q,o: [B][Sq][Hq][D] fp32; k,v: fp16, element (b,s,h,d) at
 * b*kv_bstride + s*Hkv*D + h*D + d (the cache's BSHD layout, op/CPUKVCache.cpp:10-66).
 * causal: key c visible to query r iff c <= r + (Sk - Sq).
 * ws: scratch of at least mllmcu_flash_attn_workspace_bytes(B,Sq,Hq,D). */
size_t mllmcu_flash_attn_workspace_bytes(int B, int Sq, int Hq, int D);
int mllmcu_flash_attn_f16kv(const float *q, const void *k, const void *v, float *o, int B, int Sq, int Sk, int Hq,
                            int Hkv, int D, int causal, long long kv_bstride, void *ws, size_t ws_bytes, void *stream);

/* ---- quantize_row_q8_K (ggml/QuantizeQ8.cpp:216-266): rows x k/256 native block_q8_K (292 B: float d, 256 int8,
 * 16 int16 group sums).  Bytes identical to the reference's (an all-zero super-block gets zero group sums, which the
 * reference leaves unwritten).  x rows 16-byte aligned, ldx % 4 == 0. */
int mllmcu_quantize_q8_K(const float *x, void *y_blocks, int rows, int k, int ldx, void *stream);
/* the two exact integers vec_dot_q4_K_q8_K forms per super-block (ggml/VecDotQ4.cpp:220-283):
 * sumi[n][m][i] = sum_sb scale_sb * (q4_sb . q8_sb), summ[n][m][i] = sum_sb min_sb * (bsums[2sb] + bsums[2sb+1]).
 * Test/diagnostic entry point. */
int mllmcu_block_isums_q4_K_q8_K(const void *W, const void *xq_blocks, int32_t *sumi, int32_t *summ, int M, int N, int K, void *stream);

/* ---- single-token step of a Q4_K body (LLaVA-1.5-7B): the ops in front of a Linear folded into the activation quantiser, the residual
 * add folded into the GEMV, every launch a programmatic dependent of the previous one.
 * mllmcu_prep_q8_K writes the block_q8_K rows (as mllmcu_quantize_q8_K) of
 *   mode 0: x;   mode 1: RMSNorm(x) with weight aux[k] (CPURMSNorm::execute, op/CPURMSNorm.cpp:31-95; the arithmetic of mllmcu_rmsnorm);
 *   mode 2: SiLU(x) * aux, aux laid out as x (CPUSiLU + CPUMul of the MLP, models/llama/modeling_llama.hpp:20-41; the arithmetic of
 *   mllmcu_silu_mul) -- values and bytes identical to running the separate ops and then mllmcu_quantize_q8_K.
 *   flags: 1 = programmatic dependent launch; 2 = each block re-pitched to 304 bytes (d at 0, the 256 int8 at 16 -- 16-byte aligned, which
 *   the native 292-byte block cannot be --, the 16 group sums at 272; bytes 4..15 unused; y_blocks 16-byte aligned).
 * mllmcu_gemv_q4k: y[s][m * ldy + n] = sum_k W[s][n][k] x[m][k] (+ bias[s][n]) for nW <= 3 Q4_K matrices sharing the M <= 8 activation
 * rows xq_blocks (q | k | v, gate | up: one launch) with the reference's exact integers (ggml/VecDotQ4.cpp:220-283: scale * dot and
 * min * group sum), formed per lane over 32 weights and folded into fp32 with the super-block's two factors -- only the fp32 summation
 * order differs from vec_dot_q4_K_q8_K's (as it does between its own AVX lanes).  residual != NULL (nW == 1): y = residual + that (CPUAdd
 * after o_proj / down_proj).  W, N, bias, y are HOST arrays of nW entries (bias may be NULL).  flags: 1 = programmatic dependent launch;
 * 2 = xq_blocks are the 304-byte blocks; 4 = never the variant that holds the activation row in registers (one row, K <= 4096, 304-byte
 * blocks take it by default: a persistent grid, the weight loads the only loads per row; same sums). */
int mllmcu_prep_q8_K(int mode, const float *x, const float *aux, float eps, int add_unit_offset, void *y_blocks, int rows, int k, int ldx,
                     int flags, void *stream);
int mllmcu_gemv_q4k(int nW, const void *const *W, const int *N, const float *const *bias, float *const *y, int ldy, const void *xq_blocks,
                    int M, int K, const float *residual, int ldr, int flags, void *stream);

/* ---- Q4_0_4_4 weights (block_q4_0x4, mllm/DataType.hpp:104-107; make_block_q4_0x4 + quantize_q4_0_nr_bl,
 * ggml/GemmPack.cpp:45-60, :308-344; consumed by gemv/gemm_q4_0_4x4_q8_0 when W.dtype == MLLM_TYPE_Q4_0_4_4,
 * compute/Matmul.cpp:158-217): rewritten ONCE at load into plain block_q4_0 rows (same nibbles, same scales), after which
 * the tensor is an ordinary MLLMCU_TYPE_Q4_0 weight.  nrow % 4 == 0; packed != rows. */
int mllmcu_q4_0_4x4_to_q4_0(const void *packed, void *rows, int nrow, int k, void *stream);

/* ---- GELU / QuickGELU through the reference's fp16 table: mllm_vec_gelu_f32 / mllm_vec_gelu_quick_f32
 * (ggml/Quantize.hpp:101-133): y = fp32(table[fp16 bits of x]); table_f16 = the 65536-entry table built as
 * init_table_gelu_f16 / init_table_gelu_quick_f16 build it (host, once), resident on the device. */
int mllmcu_gelu_table(const float *x, float *y, size_t n, const void *table_f16, void *stream);

/* ---- Softmax over the last dim with the causal mask of the eager attention spelling: CPUSoftMax::execute
 * (op/CPUSoftMax.cpp:29-66) + mllm_vec_soft_max_f32 (compute/ActivationFunction.cpp:29-80).  x,y: [outer][S][n];
 * row s sees (causal && S > 1) ? s + 1 + old_dim : n entries, the rest of the row is written 0.  Bit-identical to
 * the AVX2 build of the reference for the vector part (same polynomial exp, same summation order). */
int mllmcu_softmax(const float *x, float *y, int outer, int S, int n, int causal, int old_dim, void *stream);
/* x / s, IEEE division (q / sqrt(D) of the eager spelling; CPUdivFunction, op/CPUBinaryFunc.hpp) */
int mllmcu_div_scalar(const float *x, float s, float *y, size_t n, void *stream);

/* ---- M-RoPE: CPUMultimodalRoPE (op/CPUMultimodalRoPE.cpp:25-331).  The per-call tables ([S][dim], sections of the
 * t/h/w axes already gathered, built on the host as multimodal_sinusoidal_position_embedding + apply_multimodal_
 * rotary_pos_emb do) are applied rotate-half with table row s: exactly mllmcu_rope(MLLMCU_ROPE_HF, ..., pos0 = 0). */
int mllmcu_rope_mrope(const float *x, void *out, int B, int S, int H, int D, int rot_dim, const float *sin_t, const float *cos_t,
                      int tdim, int out_f16, long long out_bstride, long long out_sstride, void *stream);
/* ---- ViT 2-D rotary: CPUVisionRoPEFuncFunction::rope_hf (op/CPUVisionRoPEFunc.hpp:27-90).  x,out [B][S][H][D];
 * cos_or_null == NULL: angles_or_sin holds rotary_pos_emb [S][D/2] and sin/cos are evaluated on the device (<= 2 ulp from
 * libm's); otherwise the two pointers hold host-built sin and cos [S][D/2] and the result is bit-identical. */
int mllmcu_rope_vision(const float *x, const float *angles_or_sin, const float *cos_or_null, void *out, int B, int S, int H, int D,
                       int out_f16, void *stream);

/* ---- strided 4-D copy, the device form of the reference's physical data movers (CPUtransposeFunction
 * op/CPUTransposeFunc.hpp:127-149, CPUcatFunction, CPUsplitFunction, KV head replication op/CPUKVCache.cpp:137-247):
 * dst[sum i_j*ds[j]] = src[sum i_j*ss[j]] for i in n[0] x n[1] x n[2] x n[3]; strides in elements of elem_bytes (2 | 4).
 * n, ss, ds are HOST arrays of 4. */
int mllmcu_copy_strided(const void *src, void *dst, int elem_bytes, const int *n, const long long *ss, const long long *ds, void *stream);

/* ---- the index functions that splice image embeddings into the text prefill (models/llava/modeling_llava.hpp:128-131,
 * models/qwen2_vl/modeling_qwen2_vl.hpp:386-389).  Tensor::where (CPUwhereFunction, op/CPUWhereFunc.hpp:32-95): positions with
 * x == value of a dense BSHD tensor of `total` elements, compacted in the reference's visit order (b, s, h, d = memory order) as
 * linear indices into lin_idx (capacity total), *count on the device; mllmcu_where_unravel turns them into the coordinate of one
 * axis (0 b, 1 s, 2 h, 3 d; n = sizes in that order) as floats, which is what the reference returns. */
int mllmcu_where_eq(const float *x, long long total, float value, int *lin_idx, int *count, void *stream);
int mllmcu_where_unravel(const int *lin_idx, int count, const int *n, int axis, float *out, void *stream);
/* Tensor::index_put (CPUIndexPutFunction, op/CPUIndexPutFunc.hpp:73-131), indices read on the device: rows -- dst row (int)idx[r]
 * <- src row r (accumulate == false); splice -- out [S-1+R][dim] = dest[0:p] ++ src[0:R] ++ dest[p+1:S], p = (int)idx[0]
 * (accumulate == true with one replaced position, the LLaVA image splice). */
int mllmcu_index_put_rows(float *dst, long long dst_ld, const float *src, long long src_ld, const float *idx, int n_rows, int dim, void *stream);
int mllmcu_index_splice(const float *dest, const float *src, const float *idx, float *out, int S, int R, int dim, void *stream);
/* Tensor::range (CPURangeFunction, op/CPURangeFunc.hpp:33-41): out[i] = start + i */
int mllmcu_range(float *out, int start, int n, void *stream);

/* ---- fp32 GEMM C[b][m][n] = sum_k A[b][m*lda + k] * B[b][n*ldbn + k*ldbk] + bias[n]  (batch strides in elements):
 * the fp32 mat_mul of the eager attention spelling (CPUMatmul::execute, op/CPUMatmul.cpp:63-83: Q.K^T with ldbn = row
 * stride of K, ldbk = 1; P.V with ldbn = 1, ldbk = row stride of V) and Linears whose weights stay F32. */
int mllmcu_sgemm(const float *A, long long lda, long long strideA, const float *B, long long ldbn, long long ldbk, long long strideB,
                 const float *bias, float *C, long long ldc, long long strideC, int M, int N, int K, int batch, void *stream);
/* ---- patch-embed convolutions, VALID padding: conv2d_fp32_VALID / conv3d_fp32_VALID (compute/Convolution.cpp:35-90,
 * :179-240) as an implicit GEMM (receptive fields gathered while loading tiles).  x [B][C][H][W] (mllm BSHD with
 * S = channel), w [O][C][kh][kw], y [B][O][oh][ow], or with channels_last [B][oh*ow][O] -- the flatten(2).transpose the ViT
 * graphs apply next.  3-D: x [B][C][T][H][W], w [O][C][kt][kh][kw], y BCTHW or (channels_last) BTHWC. */
int mllmcu_conv2d_valid(const float *x, const float *w, const float *bias, float *y, int B, int C, int H, int W, int O, int kh, int kw,
                        int sh, int sw, int channels_last, void *stream);
int mllmcu_conv3d_valid(const float *x, const float *w, const float *bias, float *y, int B, int C, int T, int H, int W, int O, int kt,
                        int kh, int kw, int st, int sh, int sw, int channels_last, void *stream);

/* ---- flash attention with fp32 K/V (no cache): the vision towers' attention (flash_attention_2_forward fp32 variant,
 * compute/FlashAttention2.hpp:87-1210; models/vit/modeling_vit.hpp:42-45).  D in {32, 64, 80, 128}.  Workspace:
 * mllmcu_flash_attn_f32kv_workspace_bytes gives room for K and V split once into fp16 hi / lo planes (the tensor-core kernel then streams
 * them like a cache); with only mllmcu_flash_attn_workspace_bytes every CTA splits the tiles it reads (same values, slower on long
 * sequences). */
size_t mllmcu_flash_attn_f32kv_workspace_bytes(int B, int Sq, int Sk, int Hq, int Hkv, int D);
int mllmcu_flash_attn_f32kv(const float *q, const float *k, const float *v, float *o, int B, int Sq, int Sk, int Hq, int Hkv, int D,
                            int causal, long long kv_bstride, void *ws, size_t ws_bytes, void *stream);

/* ---- fp32 -> fp16 RNE rows (KV-cache append of a tensor that is not produced
 * by a Linear/RoPE writing in place). dst row stride ld_dst, src row stride ld_src (elements) */
int mllmcu_f32_to_f16_rows(const float *src, void *dst, int rows, int cols, long long ld_src, long long ld_dst, void *stream);
/* the reverse, exact widening: Backend::convert_fp_data (mllm/Backend.hpp:73, CPUBackend.cpp:260-277) for Tensor::fp32() */
int mllmcu_f16_to_f32_rows(const void *src, float *dst, int rows, int cols, long long ld_src, long long ld_dst, void *stream);

/* ---- optional tensor-parallel Linear (north_star): the exchange step of the row-parallel o_proj / down_proj for the
 * decode shape, as one kernel over peer memory (p2p.cu).  No counterpart in the reference (single process, CPUBackend threads).
 * Every rank cudaMalloc's one region and publishes it through CUDA IPC (get_handle -> 64 bytes to the peers -> open); a call
 * names a "site": buf_offset = where the producing GEMV of EVERY rank wrote its partial [n] floats inside its region, flag_offset
 * = `world` 32-bit words inside the region reserved for this site.  out[i] = residual[i] + sum_r partial_r[i], ranks added in
 * order 0..world-1 on every rank (bit-identical results everywhere).  *epoch_ctr (device) must hold the same, never repeated,
 * value on all ranks for one step: mllmcu_p2p_bump once per step.  peer_bases: HOST array of `world` device pointers (own region
 * at [rank]).  world <= 8, n % 4 == 0, one CTA (a decode-sized reduction; prefill goes through NCCL).  A peer that never arrives
 * traps the kernel after 10 s instead of hanging the device.  pdl != 0: launched with programmatic dependent launch (the kernel waits
 * for its producer itself and releases its dependents at once, so the next GEMV streams weights during the exchange). */
int mllmcu_ipc_get_handle(void *dev_ptr, void *handle64);
int mllmcu_ipc_open(const void *handle64, void **dev_ptr);
int mllmcu_ipc_close(void *dev_ptr);
int mllmcu_p2p_bump(unsigned *epoch_ctr, void *stream);
int mllmcu_p2p_allreduce(int world, int rank, const void *const *peer_bases, size_t buf_offset, size_t flag_offset, const unsigned *epoch_ctr,
                         const float *residual, float *out, int n, int pdl, void *stream);

/* ---- greedy sampling: LlmTextGenerator argmax (Generate.cpp:46-53) on device;
 * first maximum wins, as std::max_element does. out: int32 per row. */
int mllmcu_argmax(const float *logits, int rows, int n, int *out, void *stream);
/* same, writing the winning index as an fp32 token id (mllm feeds token ids as fp32,
 * tokenizers/Tokenizer.hpp:78-101) so the next step's input never leaves the device */
int mllmcu_argmax_ids(const float *logits, int rows, int n, float *ids_out, void *stream);

/* ---- top-k / top-p sampling on the device: _LlmTextGenerateTopkSamplingMethod::generate (Generate.cpp:45-91, mode 0, p_or_k = k in
 * 2..1024) and _LlmTextGenerateToppSamplingMethod::generate (Generate.cpp:93-143, mode 1, p_or_k = p; x must hold probabilities,
 * at most 1024 candidates).  x [rows][n] (row stride ldx); candidates in descending order, softmax with temperature in the
 * reference's float / double mix, renormalised, and the draw of std::discrete_distribution with the caller's u[row] in [0, 1)
 * in place of the reference's unseeded generator.  ws: rows * n floats of scratch; ids_out[row] = the sampled id as fp32; optional
 * cand_p / cand_idx [rows][cand_cap] and n_cand [rows] return the candidate distribution (what the parity tests compare); n_cand < 0:
 * top-p hit the 1024-candidate limit before its running sum reached p (the draw is from those 1024). */
int mllmcu_sample(int mode, const float *x, int rows, int n, long long ldx, float p_or_k, float temperature, const float *u, float *ws, float *ids_out,
                  float *cand_p, int *cand_idx, int *n_cand, int cand_cap, void *stream);

/* ---- single-token decode step, graph-replayable: per-step scalars live in device memory.
 * ctl[0] = position of the token being consumed (== KV length before its append).
 * mllmcu_decode_embed: CPUEmbedding Q4_0 gather (op/CPUEmbedding.cpp:66-81) and ctl[0] += 1 (ctl may be NULL). */
int mllmcu_decode_embed(const void *table, const float *ids, float *out, int n_ids, int hidden, int *ctl, void *stream);
/* q_rope + k_rope (op/CPURoPE.cpp:200-372) on the fused q|k|v projection rows qkv[B][ld] (q rotated in place)
 * and k_cache / v_cache append in fp16 at cache row ctl[0] (op/CPUKVCache.cpp:137-247). */
int mllmcu_decode_rope_kv(int rope_type, float *qkv, int ld, int B, int Hq, int Hkv, int D, int rot_dim, const float *sin_t,
                          const float *cos_t, int tdim, const int *ctl, void *kcache, void *vcache, long long kv_bstride,
                          int cache_limit, int pdl, void *stream);
/* flash_attention_2_forward, decode (Sq == 1) over ctl[0]+1 cached keys (compute/FlashAttention2.hpp:1212-2284),
 * split over keys with the merge done by the last CTA to finish.  D in {32, 64, 128}.  ws must be zeroed once. */
int mllmcu_decode_attn_nsplit(int B, int Hq);
size_t mllmcu_decode_attn_workspace_bytes(int B, int Hq, int D);
int mllmcu_decode_attn(const float *q, int ldq, const void *k, const void *v, float *o, int ldo, const int *ctl, int B, int Hq,
                       int Hkv, int D, long long kv_bstride, void *ws, size_t ws_bytes, int pdl, void *stream);
/* the same with q_rope / k_rope / k_cache / v_cache fused in (full rotary only: rot_dim == D == tdim): qkv holds the raw
 * q|k|v projection rows and is not modified; k (rotated) and v are appended in fp16 at cache row ctl[0] */
int mllmcu_decode_attn_rope(int rope_type, const float *qkv, int ld, void *kcache, void *vcache, float *o, int ldo, const int *ctl,
                            int B, int Hq, int Hkv, int D, const float *sin_t, const float *cos_t, int tdim, long long kv_bstride,
                            int cache_limit, void *ws, size_t ws_bytes, int pdl, void *stream);
/* greedy pick (Generate.cpp:46-53): ids_out[b] fp32 and, when tok_log != NULL, tok_log[ctl[0]*rows + b] (ctl[0] < log_rows) */
int mllmcu_decode_argmax(const float *logits, int rows, int n, float *ids_out, int *tok_log, const int *ctl, int log_rows,
                         void *stream);

/* ---- the whole decode step as ONE persistent kernel (decode_step.cu).  The caller describes the step as a list
 * of fused Linears ("jobs", same fusion vocabulary as mllmcu_gemv_fused: optional RMSNorm prologue, epilogue
 * 0 plain / 1 + residual / 2 silu(gate)*up); a job with post = 1 is followed by RoPE + KV append + decode
 * attention of block `layer` (mllmcu_decode_rope_kv / mllmcu_decode_attn semantics, reading the job's output
 * q|k|v rows in desc.qkv and writing desc.ao).  Weights stream through per-warp TMA rings that run ahead across
 * job boundaries; activations cross jobs through a grid barrier.  create() copies the tables to the device
 * (cudaMalloc: the plan owns ~100 KB) and returns MLLMCU_ENOSUP when a shape does not fit the kernel (the
 * caller then chains the per-op kernels).  Replaces, per token, the reference's whole op sequence of
 * LLaMAModel::Forward (models/llama/modeling_llama.hpp:98-115) for S == 1. */
typedef struct mllmcu_step_job {
    const void *W[3];
    const float *bias[3];
    int N[3];
    int nW;
    int K;
    int epi;
    int post;
    int layer;
    const float *x;
    int ldx;
    const float *norm_w;
    float eps;
    const float *residual;
    int ldr;
    float *y;
    int ldy;
} mllmcu_step_job;
typedef struct mllmcu_step_desc {
    int B, Hq, Hkv, D, rope_type, rot_dim, tdim, cache_limit, n_layers;
    const float *sin_t, *cos_t;
    long long kv_bstride;
    void *const *kcache; /* host array [n_layers] of device pointers */
    void *const *vcache;
    float *qkv;
    int ldqkv;
    float *ao;
    int ldao;
    const int *ctl;
} mllmcu_step_desc;
int mllmcu_decode_step_create(const mllmcu_step_desc *desc, const mllmcu_step_job *jobs, int njobs, void **plan);
int mllmcu_decode_step_launch(void *plan, void *stream);
int mllmcu_decode_step_destroy(void *plan);

#ifdef __cplusplus
}
#endif
#endif
