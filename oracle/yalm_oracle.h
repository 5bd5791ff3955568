/* TEST INFRASTRUCTURE — NOT PRODUCT CODE.
 *
 * CPU restatement, in plain C, of the reference's token-generation forward pass
 * (xiaguan/yalm src/infer.cpp).  Only tests/, __graft_entry__.smoke() and bench.py's
 * cpu_baseline / --impl reference legs may use it, and only as the checker.
 *
 * Parity status: PINNED.  tests/test_oracle.py checks this restatement against
 *   (1) the reference's only known-answer vector, test_attn (src/test.cpp:68-126), and
 *   (2) outputs of the unmodified reference built into oracle/_ref (logits / KV / greedy tokens of
 *       seeded tiny models, committed under tests/golden/ with the generating script).
 *
 * Beyond the reference it adds one thing the reference has no arithmetic for:
 * fp8 e4m3 weights with one fp32 scale per output row, computed as the reference fp32 path run on
 * the de-quantised weights.
 */
#ifndef YALM_ORACLE_H
#define YALM_ORACLE_H
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

enum { YO_F32 = 0, YO_F16 = 1, YO_F8E4M3 = 4 }; /* values of the reference DType enum, src/codec.h:15-25 */
enum { YO_GELU = 0, YO_SILU = 1 };              /* ActivationType, src/model.h:14-17 */

typedef struct {
	int dim, hidden_dim, head_dim, n_layers, n_heads, n_kv_heads, vocab_size, max_seq_len;
	float rope_theta;
	int rotary_dim;
	float norm_eps;
	int act;
	float qkv_clip;
	int n_experts, n_experts_active;
	int weight_dtype;
} yo_config;

typedef struct {
	const float *rms_att, *rms_ffn;
	const void *wq, *wk, *wv, *wo, *w1, *w2, *w3, *moegate;
	/* per-row scales, fp8 only (NULL otherwise) */
	const float *sq, *sk, *sv, *so, *s1, *s2, *s3, *sgate;
	uint16_t *key_cache, *value_cache; /* (max_seq_len, n_kv_heads*head_dim) fp16, owned by caller */
} yo_layer;

typedef struct {
	yo_config c;
	const void *embed;
	const float *embed_scale;
	const float *rms_final;
	const void *wcls;
	const float *wcls_scale;
	yo_layer *layers;
	/* scratch, all caller-allocated: see yo_state_floats() */
	float *x, *xb, *xb2, *hb, *hb2, *q, *k, *v, *att, *logits, *moe_weights, *active_w;
	int *active_e;
} yo_model;

/* y = W x for a (d,n) row-major W; scale may be NULL. src/infer.cpp:48-98 */
void yo_matmul(float *xout, const float *x, const void *w, const float *scale, int n, int d, int dtype);
void yo_rmsnorm(float *o, const float *x, const float *weight, int size, float eps);            /* :134-144 */
void yo_rope(float *vec, int d, int head_dim, int pos, float theta, int rotary_dim);              /* :200-213 */
void yo_attn(float *xout, float *atth, const float *qh, const uint16_t *kh, const uint16_t *vh, int head_dim,
             int n_kv_heads, int kv_len);                                                         /* :216-248 */
void yo_mha(float *xout, float *att, const uint16_t *kb, const uint16_t *vb, const float *q, int head_dim,
            int kv_len, int max_seq_len, int n_heads, int n_kv_heads);                            /* :387-403 */
void yo_moe_gate(float *moe_weights, int *active_experts, const float *x, int n_experts, int n_active); /* :100-132 */
void yo_ffn(float *xout, const float *x, const void *w1, const void *w2, const void *w3, const float *s1,
            const float *s2, const float *s3, int hidden_dim, int dim, int act, int dtype);       /* :412-438 */
void yo_block(yo_model *m, int layer, int pos, int kv_sink, int kv_pos, int kv_len);              /* :254-385 */
/* mode 0 = HYDRATE_KV_CACHE, 1 = OUTPUT_LOGITS. src/infer.cpp:474-524 */
void yo_forward(yo_model *m, int token, int pos, int mode);
int yo_argmax(const float *logits, int n);                                                        /* src/sampler.cpp:27-38 */
float yo_half_to_float(uint16_t h);
uint16_t yo_float_to_half(float f);
float yo_e4m3_to_float(uint8_t b);

#ifdef __cplusplus
}
#endif
#endif
