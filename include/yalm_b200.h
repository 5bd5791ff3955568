/* yalm_b200 — C ABI of the B200-native token-generation forward pass.
 *
 * This is the drop-in boundary for xiaguan/yalm's `-d cuda` path.  Three layers:
 *
 *  (1) the seven memory/stream helpers the reference already exports from src/infer.cu and declares
 *      in src/model.h:33-39 — same names, same meaning;
 *  (2) one entry point per fused device op (device pointers + sizes + stream, returns the
 *      cudaError_t as int) — what replaces the kernel launches inside Block::_block_cuda
 *      (src/infer.cu:699-888) and Model::_forward_cuda_build_graph (src/infer.cu:1041-1128);
 *  (3) host-pointer test entry points with the semantics of the reference's "exposed for tests"
 *      functions matmul_cuda / mha_cuda / ffn_cuda (src/model.h:353-384, src/infer.cu:890-1007),
 *      and opaque-handle wrappers over Model / InferenceState (src/model.h:84-327) for bindings
 *      that cannot consume C++ classes (ctypes, cgo, JNI).
 *
 * Error convention: layer (1) keeps the reference's (CUDA failure -> message on stderr + abort(),
 * src/infer.cu:13-31).  Layers (2)/(3) return 0 on success, a cudaError_t or a negative YALM_E*
 * code otherwise, and never fall back to a CPU path.
 */
#ifndef YALM_B200_H
#define YALM_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* Weight element types; values are the reference's DType enum (src/codec.h:15-25). */
#define YALM_F32 0
#define YALM_F16 1
#define YALM_F8E5M2 3 /* parsed, but no arithmetic exists for it: rejected */
#define YALM_F8E4M3 4 /* net-new: e4m3 codes + one fp32 scale per output row */

#define YALM_ACT_GELU 0 /* ActivationType, src/model.h:14-17 */
#define YALM_ACT_SILU 1

#define YALM_MODE_HYDRATE 0 /* InferenceMode, src/model.h:28-31 */
#define YALM_MODE_LOGITS 1

#define YALM_EINVAL (-1)  /* unsupported shape / dtype / null pointer */
#define YALM_ENOGPU (-2)  /* no CUDA device: there is deliberately no CPU fallback */
#define YALM_EFORMAT (-3) /* bad .yalm container */

typedef void *yalm_stream_t; /* cudaStream_t */

/* ---- (1) reference C ABI, src/model.h:33-39 / src/infer.cu:59-90 ------------------------------ */
void *upload_cuda(void *host, size_t size);            /* cudaMalloc + H2D, returns device pointer */
void *download_cuda_raw(void *device, size_t size);    /* C-clean form of download_cuda (the reference's takes a std::string) */
void register_cuda_host(void *host, size_t size);
void free_cuda(void *device);
void unregister_cuda_host(void *host);
void set_cuda_device(int device);
void init_cuda_stream(yalm_stream_t *stream);

/* ---- (2) fused device ops ---------------------------------------------------------------------
 * Decode control block `ctl` (device, 8 ints): [0] token [1] pos [2] kv_sink [3] kv_pos [4] kv_len
 * [5] step counter [6..7] reserved.  Keeping pos/kv_* in device memory is what lets one CUDA graph be
 * replayed for every token with zero per-token node updates (the reference patches 129-161 kernel
 * nodes per token, src/infer.cu:1146-1163).                                                        */
#define YALM_CTL_INTS 8

/* x = float(E[token,:]) (src/infer.cu:622-640) and derive the KV ring position from pos
 * (src/infer.cu:1081-1083: kv_sink/kv_pos/kv_len with KV_SINKS = 2) into ctl.  token >= 0: token and
 * pos are taken from the arguments (copied at launch, so back-to-back calls never race); token < 0:
 * ctl[0..1] are taken as already set on the device (by yalm_argmax_advance). */
int yalm_embed(float *x, const void *table, const float *row_scale, int dim, int dtype, int *ctl, int token,
               int pos, int max_seq_len, yalm_stream_t s);

/* KV cache layout on the device: HEAD-MAJOR fp16 ring, (n_kv_heads, max_seq_len, head_dim) — one contiguous
 * slab per KV head, so a run of rows of one head is one 1-D bulk (TMA) copy.  The reference keeps
 * (max_seq_len, n_kv_heads, head_dim) (src/model.cpp:171-172); the host-pointer test entry yalm_mha_host and
 * the yalm_state_read taps take / return the reference layout and convert. */

/* rmsnorm -> q,k,v GEMVs -> clip -> RoPE(q,k) -> fp16 KV append at ctl.kv_pos -> sink re-rotation.
 * Replaces rmsnorm + fused_qkv_matmul_clip + fused_rope_and_cache_update + rotate_sink_tokens
 * (src/infer.cu:526-539, 291-336, 642-697).  Weights are (out,in) row-major, y = W x. */
int yalm_qkv_rope_append(float *q_out, uint16_t *key_cache, uint16_t *value_cache, const float *x,
                         const float *rms_weight, const void *wq, const void *wk, const void *wv,
                         const float *sq, const float *sk, const float *sv, int dim, int n_heads, int n_kv_heads,
                         int head_dim, int rotary_dim, int max_seq_len, float rope_theta, float norm_eps,
                         float qkv_clip, int dtype, const int *ctl, yalm_stream_t s);

/* Split-K flash-decode GQA attention over the fp16 KV ring: replaces attn_dot + attn_softmax +
 * att_mix (src/infer.cu:338-524).  `workspace` holds the per-split partials; size from
 * yalm_attn_workspace_bytes().  out is (n_heads*head_dim) fp32. */
size_t yalm_attn_workspace_bytes(int n_heads, int n_kv_heads, int head_dim);
int yalm_attn_decode(float *out, const uint16_t *key_cache, const uint16_t *value_cache, const float *q,
                     void *workspace, int n_heads, int n_kv_heads, int head_dim, int max_seq_len, const int *ctl,
                     yalm_stream_t s);

/* yalm_attn_decode without a caller workspace and without any allocation (no cudaMalloc / cudaFree / allocating library
 * call).  This entry point has no counterpart in the reference: it was added for
 * callers that cannot allocate (code under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).
 * The split-K partials live in a file-scope device array of yalm_attn_decode_static_max_bytes() = 2 MiB that every GPU
 * loading the library holds until exit; shapes whose yalm_attn_workspace_bytes() exceeds it (or with more than 64 KV
 * heads) are refused with YALM_EINVAL and a message on stderr.  All calls on one device share that array: calls that
 * overlap on the GPU (different streams or host threads) race — serialise them on one stream.  Same result as
 * yalm_attn_decode (same kernel, same split count). */
size_t yalm_attn_decode_static_max_bytes(void);
int yalm_attn_decode_static(float *out, const uint16_t *key_cache, const uint16_t *value_cache, const float *q, int n_heads,
                            int n_kv_heads, int head_dim, int max_seq_len, const int *ctl, yalm_stream_t s);

/* out[i] = (residual ? residual[i] : 0) + sum_k aw[k] * (W[e_k] v_k)[i]   (wo and w2 + residual:
 * fused_matmul_add_residuals, src/infer.cu:270-288; MoE mix: src/infer.cpp:377-383).
 * Dense: n_active = 1, active_e = active_w = NULL.  MoE: expert e's matrix starts at
 * w + e*d*n elements, its scales at scale + e*d, and v_k = v + k*n. */
int yalm_gemv_residual(float *out, const float *residual, const float *v, const void *w, const float *scale, int n,
                       int d, int dtype, int n_active, const int *active_e, const float *active_w,
                       yalm_stream_t s);

/* Expert-parallel variants (the measured alternative to slicing every expert): the rank holds only the
 * experts e with e % ep_size == ep_rank, stacked in ascending order in w / w1 / w3 (and their scales); the router is
 * replicated and still selects among all n_experts; rows of experts held elsewhere are skipped here and reach x through
 * the all-reduce that follows w2.  ep_size == 1 is the plain function. */
int yalm_gemv_residual_ep(float *out, const float *residual, const float *v, const void *w, const float *scale, int n, int d,
                          int dtype, int n_active, const int *active_e, const float *active_w, int ep_rank, int ep_size,
                          yalm_stream_t s);
int yalm_ffn_glu_ep(float *hb, const float *x, const float *rms_weight, const void *w1, const void *w3, const float *s1,
                    const float *s3, const void *gate, const float *sgate, int dim, int hidden_dim, int act, float norm_eps,
                    int dtype, int n_experts, int n_active, int *active_e, float *active_w, float *moe_logits, int ep_rank,
                    int ep_size, yalm_stream_t s);

/* rmsnorm -> [router GEMV + top-k gate] -> hb[k,i] = act(W1[e_k] xb)[i] * (W3[e_k] xb)[i].
 * Replaces rmsnorm + fused_ffn_w1_w3_glu_act (src/infer.cu:598-620) and, for MoE, the CPU-only
 * router (src/infer.cpp:100-132, 347-349).  Dense: n_experts = 0, gate = NULL; hb is (hidden).
 * MoE: hb is (n_active, hidden); active_e/active_w (device, n_active) receive the selection.
 * rms_weight == NULL skips the norm (the reference's ffn_cuda test entry has none). */
int yalm_ffn_glu(float *hb, const float *x, const float *rms_weight, const void *w1, const void *w3,
                 const float *s1, const float *s3, const void *gate, const float *sgate, int dim, int hidden_dim,
                 int act, float norm_eps, int dtype, int n_experts, int n_active, int *active_e,
                 float *active_w, float *moe_logits, yalm_stream_t s);

/* logits = Wcls rmsnorm(x)  (final rmsnorm + matmul_wide, src/infer.cu:1097-1126).  logits may be
 * device memory or device-mapped host memory. */
int yalm_lm_head(float *logits, const float *x, const float *rms_weight, const void *wcls, const float *scale,
                 int dim, int vocab, float norm_eps, int dtype, yalm_stream_t s);

/* Greedy on device with Sampler::sample_argmax's first-max rule (src/sampler.cpp:27-38): writes the
 * arg-max to tokens_out[ctl.step], then ctl.token = argmax, ctl.pos += 1, ring position re-derived. */
int yalm_argmax_advance(const float *logits, int vocab, int *ctl, int max_seq_len, int *tokens_out,
                        yalm_stream_t s);

/* Sampler::sample on the device (src/sampler.cpp:40-65): p_i = exp((l_i - max) / T) / sum, the first index whose
 * running sum reaches r wins (vocab - 1 if none); r = uniform(seed, ctl.step) from a counter-based generator (the
 * reference draws std::rand()).  Then the same control-block advance as yalm_argmax_advance.  temperature 0 = greedy. */
int yalm_sample_advance(const float *logits, int vocab, int *ctl, int max_seq_len, int *tokens_out, float temperature,
                        unsigned long long seed, yalm_stream_t s);
/* One draw into *token_dev (device int) with r = uniform(seed, counter); temperature 0 = first-max arg-max. */
int yalm_sample_device1(const float *logits, int vocab, float temperature, unsigned long long seed,
                        unsigned long long counter, int *token_dev, yalm_stream_t s);
/* Sampler::sample_prob (src/sampler.cpp:11-25) for n_rows logit rows (device, row stride vocab):
 * probs[i] = softmax(logits[i, :])[targets[i]] — what perplexity mode needs per position (src/main.cpp:174-184). */
int yalm_token_probs(const float *logits, int vocab, int n_rows, const int *targets, float *probs, yalm_stream_t s);

/* Bare GEMV xout = W x (matmul<T>, src/infer.cu:234-248), device pointers. */
int yalm_gemv(float *xout, const float *x, const void *w, const float *scale, int n, int d, int dtype,
              yalm_stream_t s);

/* ---- (2b) batched prefill ops (net-new, SURVEY.md 8(f).1: the reference hydrates and scores token by token,
 * src/main.cpp:91-97, 174-184) ---------------------------------------------------------------------------------------
 * c (M, ldc) fp32 (=|+=) a (M, K) fp16 * w (N, K)^T fp16 on the tcgen05 tensor cores (TMA operand loads, TMEM
 * accumulator); w is a .yalm weight as stored ((out, in) row-major).  accumulate != 0: c += ... (residual in place). */
int yalm_gemm_f16(float *c, int ldc, const void *a, const void *w, int M, int N, int K, int accumulate, yalm_stream_t s);
/* h (M, ldh) fp16 = act(a w1^T) * (a w3^T): fused w1 / w3 projection + GLU (src/infer.cpp:360-375). */
int yalm_gemm_glu_f16(void *h, int ldh, const void *a, const void *w1, const void *w3, int M, int N, int K, int act,
                      yalm_stream_t s);
/* Causal attention of n query rows q16 (n, n_heads * head_dim) fp16 = positions pos0 .. pos0 + n - 1 over the head-major
 * KV ring (rows 0 .. pos0 + n - 1 valid; no wrap: pos0 + n <= max_seq_len); out16 like q16.  src/infer.cpp:216-248 per row. */
int yalm_attn_prefill(void *out16, const void *q16, const uint16_t *key_cache, const uint16_t *value_cache, int n, int pos0,
                      int n_heads, int n_kv_heads, int head_dim, int max_seq_len, yalm_stream_t s);

/* ---- (3a) host-pointer test entry points (src/infer.cu:890-1007 semantics) ------------------- */
int yalm_matmul_host(float *xout, const float *x, const void *w, const float *scale, int n, int d, int dtype);
int yalm_mha_host(float *xout, float *att, const uint16_t *kb, const uint16_t *vb, const float *q, int head_dim,
                  int kv_len, int max_seq_len, int n_heads, int n_kv_heads);
int yalm_ffn_host(float *xout, const float *x, const void *w1, const void *w2, const void *w3, const float *s1,
                  const float *s2, const float *s3, int hidden_dim, int dim, int act, int dtype);
/* Host-pointer test entries of the device sampler: one draw with a caller-supplied r in [0, 1] (Sampler::sample with
 * std::rand() / RAND_MAX replaced by r), and sample_prob for n_rows rows. */
int yalm_sample_host(const float *logits, int vocab, float temperature, float r, int *token);
int yalm_token_probs_host(const float *logits, int vocab, int n_rows, const int *targets, float *probs);
/* Host-pointer test entry of the prefill GEMM: c (M, N) fp32 = a (M, K) fp16 * w (N, K)^T fp16 [+ c if accumulate];
 * with w3 != NULL instead h (M, N) fp16 = act(a w^T) * (a w3^T). */
int yalm_gemm_host(float *c, void *h, const void *a, const void *w, const void *w3, int M, int N, int K, int accumulate, int act);
/* Same ops timed on the device: `iters` launches after `warmup`, returns mean microseconds per
 * launch (CUDA events) in *usec.  This is what `./build/test -k <name>` prints. */
int yalm_bench_op(const char *name, int iters, int warmup, double *usec, double *bytes);

/* ---- (3b) opaque handles over Model / InferenceState ------------------------------------------ */
typedef struct yalm_model yalm_model;
typedef struct yalm_state yalm_state;

/* `context` = the reference's -T (0: min(metadata max_seq_len, 4096), src/model.cpp:31-36). */
int yalm_model_open_file(const char *path, int context, yalm_model **out);
/* Image must stay valid until yalm_model_cuda() has returned. */
int yalm_model_open_memory(const void *image, size_t size, int context, yalm_model **out);
/* Tensor-parallel shard `tp_rank` of `tp_size` (1 = whole model) is uploaded to `device`. */
int yalm_model_cuda(yalm_model *m, int device, int tp_rank, int tp_size);
void yalm_model_close(yalm_model *m);
/* out[16]: dim, hidden_dim, head_dim, n_layers, n_heads, n_kv_heads, vocab_size, max_seq_len,
 * rotary_dim, n_experts, n_experts_active, weight dtype, act, tp_rank, tp_size, reserved */
void yalm_model_config(const yalm_model *m, int *out);
size_t yalm_model_device_bytes(const yalm_model *m);
/* Loader facts (the reference: one cudaMalloc + one pageable copy per tensor, src/model.cpp:185-211):
 * out[0] = device allocations behind the model's weights and KV caches (1: one arena per rank),
 * out[1] = 1 if the upload went through the pinned double-buffered staging stream, 0 if it fell back to pageable copies. */
void yalm_model_load_info(const yalm_model *m, int *out2);

int yalm_state_new(yalm_model *m, yalm_state **out);
void yalm_state_free(yalm_state *s);
/* Model::forward (src/model.cpp:396-407). After a YALM_MODE_LOGITS call returns, yalm_logits() is valid. */
int yalm_forward(yalm_model *m, yalm_state *s, int token, int pos, int mode);
float *yalm_logits(yalm_state *s); /* host, vocab_size fp32 */
/* n_gen greedy steps entirely on the device starting from (token,pos): per step forward + argmax;
 * tokens_out (host, n_gen ints) gets the generated ids.  If logits_out != NULL (host, n_gen*vocab)
 * each step's logits are copied back too (parity mode).  *device_ms (optional) = CUDA-event time. */
int yalm_generate_greedy(yalm_model *m, yalm_state *s, int token, int pos, int n_gen, int *tokens_out,
                         float *logits_out, float *device_ms);
/* Batched prefill / perplexity: positions pos0 .. pos0 + n - 1 in one pass (dense tcgen05 GEMMs over the chunk, causal
 * flash attention, bulk KV append).  Afterwards the KV cache holds the chunk and yalm_logits() the LAST position's
 * logits; if targets != NULL, probs_out[i] = softmax(logits_i)[targets[i]] (Sampler::sample_prob per position,
 * src/sampler.cpp:11-25).  n <= yalm_model_batch_capacity() (0: this model / shard has no batched path — fp16 dense
 * unsharded models have — and the caller walks the tokens with yalm_forward); pos0 + n <= max_seq_len (no ring wrap). */
int yalm_model_batch_capacity(const yalm_model *m);
int yalm_forward_batch(yalm_model *m, yalm_state *s, const int *tokens, int n, int pos0, const int *targets, float *probs_out);
/* The same loop with Sampler::sample on the device: temperature 0 = greedy, > 0 = one draw per step with
 * r = uniform(seed, step).  4 bytes per token come back (the reference moves vocab * 4, src/model.cpp:338). */
int yalm_generate(yalm_model *m, yalm_state *s, int token, int pos, int n_gen, float temperature,
                  unsigned long long seed, int *tokens_out, float *logits_out, float *device_ms);
/* Tests only: set the launch counter behind the megakernel's hand-over flags (exercises its wrap). */
void yalm_state_debug_set_launch_seq(yalm_state *s, unsigned seq);
/* Debug/parity taps: copy a device activation or one layer's KV rows to the host.
 * which: 0 x, 1 xb2(attn out), 2 hb, 3 q, 4 key_cache(layer), 5 value_cache(layer),
 * 6 active_experts (int), 7 active_experts_weights */
int yalm_state_read(yalm_model *m, yalm_state *s, int which, int layer, void *dst, size_t bytes);
/* Mean duration (microseconds, CUDA events on the launching stream) of the dominant kernel — the fused
 * rmsnorm + w1/w3 + GLU GEMV — over n_steps greedy decode steps x n_layers launches from (token,pos). */
int yalm_profile_ffn(yalm_model *m, yalm_state *s, int token, int pos, int n_steps, double *mean_us);
/* Number of kernel launches one forward pass issues in `mode` (for bench.py's gpu_launches). */
int yalm_forward_launches(const yalm_model *m, int mode);
/* 1 when forward() is one persistent cooperative kernel per token (unsharded dense models; the weight and KV
 * tiles of the whole pass stream through a shared-memory TMA ring), 0 when it is the per-op kernel graph
 * (tensor-parallel shards, MoE, shapes the tile planner rejects, or YALM_B200_MEGA=0 in the environment). */
int yalm_model_uses_megakernel(const yalm_model *m);
/* Plan of the whole-token kernel for this model (no reference counterpart: the reference has no such kernel):
 * out8 = {ring stages, floats of the shared-memory vector region, 1 if hb travels w1/w3 -> w2 as pipelined slices
 * (decode_mega.cuh "pipelined hb") else 0, K splits of w2, 1 if the GEMV phases run on the tensor cores (decode_mega.cuh
 * "tensor-core GEMV") else 0, 0, 0, 0}.  Returns YALM_EINVAL when the model does not run on it. */
int yalm_model_mega_info(const yalm_model *m, int *out8);
/* One forward pass (logits mode) with the megakernel's phase profile on.  out receives grid * n_layers * 48
 * clock64 values: per CTA, per layer, 6 phases (qkv, attention, wo, w1/w3, w2, lm head) x 8 slots
 * ([0] phase entered [1] grid barrier passed [2] input vector staged [3] tiles + epilogue done
 * [4] cycles spent waiting for ring stages to fill).  out == NULL: only *grid (CTAs) is returned. */
int yalm_mega_profile(yalm_model *m, yalm_state *s, int token, int pos, long long *out, size_t count, int *grid);

/* Tensor-parallel plumbing (one process per GPU).  The 128-byte NCCL unique id is created by rank 0
 * with yalm_tp_unique_id() and distributed by the caller (torch.distributed / MPI / files). */
int yalm_tp_unique_id(void *id128);
int yalm_tp_init(const void *id128, int rank, int world);
void yalm_tp_finalize(void);
/* Shard arithmetic of the loader, callable without a device (tests): which bytes of an unsharded row-major tensor
 * rank `rank` of `size` holds.  dims[7] = dim, hidden, head_dim, n_heads, n_kv_heads, vocab, bytes per weight;
 * tensor: 0 wq, 1 wk, 2 wv, 3 wo, 4 w1, 5 w2, 6 w3, 7 wcls; out[4] = spitch, offset, width, height (bytes; `height`
 * pieces of `width`, piece i at offset + i * spitch).  Column-parallel: wq wk wv w1 w3 wcls; row-parallel: wo w2
 * (the reference is single-GPU). */
int yalm_tp_shard_plan(const int *dims, int tensor, int rank, int size, long long *out);

/* Planning arithmetic of the whole-token kernel, callable without a device (tests).  It replaces the launch geometry
 * the reference hard-codes per kernel (src/infer.cu:745-880: one warp per output row, grids of d / 32 CTAs).
 * yalm_mega_tile_plan: out20 = 5 GEMV kinds (qkv, wo, w1/w3, w2, lm head) x {rows per tile, columns per tile, K splits,
 * row groups} for one rank's dimensions; row group g of a phase goes to CTA (rotation + g) % grid.
 * yalm_mega_attn_split: out5 = {kv head, first CTA of the head, CTAs of the head, first chunk, chunks} of CTA `cta` for a
 * context of `n_chunks` KV row chunks per head.  Both return YALM_EINVAL for shapes the kernel does not take. */
int yalm_mega_tile_plan(int dtype, int dim, int q_dim, int kv_dim, int hidden, int vocab, int *out20);
/* Tensor-core GEMV geometry (decode_mega.cuh "tensor-core GEMV"): out4 = {rows per tile, first row, rows, tiles} of CTA `cta`
 * of `grid` for a matrix of `rows` rows partitioned in units of `unit` rows with tiles of at most `cap` rows.
 * yalm_mega_hb_share (pipelined hb): the number of hb elements of slice `slice` (slice_elems elements each) that CTA `cta`
 * produces when the w1/w3 phase has `groups` row groups of rows_per_group rows, group g on CTA (rotation + g) % grid;
 * negative = YALM_EINVAL. */
int yalm_mega_mma_plan(int rows, int unit, int cap, int grid, int cta, int *out4);
int yalm_mega_hb_share(int groups, int rows_per_group, int slice_elems, int grid, int cta, int rotation, int slice);
int yalm_mega_attn_split(int cta, int ncta, int n_kv_heads, int n_chunks, int *out5);

/* The CLI's tokenizer (greedy longest-prefix match + byte fallback, src/tokenizer.cpp:57-94; decode_one :44-55) over an
 * in-memory .yalm image; host only.  yalm_tokenize returns the number of ids (out receives at most cap of them),
 * yalm_detokenize the byte length of the text (out receives at most cap - 1 bytes + NUL); negative = error. */
int yalm_tokenize(const void *image, size_t size, const char *text, int add_bos, int *out, int cap);
int yalm_detokenize(const void *image, size_t size, const int *ids, int n, char *out, int cap);

const char *yalm_error_string(int code);
const char *yalm_version(void);

#ifdef __cplusplus
}
#endif
#endif /* YALM_B200_H */
