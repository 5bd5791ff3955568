/* TEST INFRASTRUCTURE — NOT PRODUCT CODE.  See yalm_oracle.h for scope and parity status.
 * Each function cites the reference lines (xiaguan/yalm) whose arithmetic it restates. */
#include "yalm_oracle.h"

#include <float.h>
#include <immintrin.h>
#include <math.h>
#include <string.h>

float yo_half_to_float(uint16_t h) { return _cvtsh_ss(h); }      /* src/infer.cpp:10-12 */
uint16_t yo_float_to_half(float f) { return _cvtss_sh(f, 0); }   /* src/infer.cpp:13-15 (round-to-nearest-even) */

float yo_e4m3_to_float(uint8_t b) {
	int e = (b >> 3) & 15, m = b & 7;
	float mag = e == 0 ? ldexpf((float)m, -9) : ldexpf(1.0f + (float)m / 8.0f, e - 7);
	if ((b & 0x7f) == 0x7f) mag = NAN;
	return (b & 0x80) ? -mag : mag;
}

/* src/infer.cpp:48-59 (fp32: one running sum per row, j ascending) and :63-98 (fp16: 16 interleaved
 * partial sums updated with fused multiply-adds, lanes 0-7 <- elements j..j+7, lanes 8-15 <- j+8..j+15,
 * then lo+hi, then fold 8->4, then a left-to-right 4-term sum as _mm_dp_ps does). */
void yo_matmul(float *xout, const float *x, const void *w, const float *scale, int n, int d, int dtype) {
	int i;
#pragma omp parallel for private(i)
	for (i = 0; i < d; i++) {
		float val = 0.0f;
		if (dtype == YO_F32) {
			const float *wr = (const float *)w + (size_t)i * n;
			for (int j = 0; j < n; j++) val += wr[j] * x[j];
		} else if (dtype == YO_F16 && n % 16 == 0) {
			const uint16_t *wr = (const uint16_t *)w + (size_t)i * n;
			float s[16];
			for (int l = 0; l < 16; l++) s[l] = 0.0f;
			for (int j = 0; j < n; j += 16)
				for (int l = 0; l < 16; l++) s[l] = fmaf(yo_half_to_float(wr[j + l]), x[j + l], s[l]);
			float s8[8], s4[4];
			for (int l = 0; l < 8; l++) s8[l] = s[l] + s[l + 8];
			for (int l = 0; l < 4; l++) s4[l] = s8[l] + s8[l + 4];
			val = ((s4[0] + s4[1]) + s4[2]) + s4[3];
		} else if (dtype == YO_F16) {
			const uint16_t *wr = (const uint16_t *)w + (size_t)i * n;
			for (int j = 0; j < n; j++) val += yo_half_to_float(wr[j]) * x[j];
		} else { /* fp8 e4m3: the fp32 loop on de-quantised weights */
			const uint8_t *wr = (const uint8_t *)w + (size_t)i * n;
			float sc = scale ? scale[i] : 1.0f;
			for (int j = 0; j < n; j++) val += (yo_e4m3_to_float(wr[j]) * sc) * x[j];
		}
		xout[i] = val;
	}
}

/* src/infer.cpp:134-144 */
void yo_rmsnorm(float *o, const float *x, const float *weight, int size, float eps) {
	float rms = 0.0f;
	for (int i = 0; i < size; ++i) rms += x[i] * x[i];
	rms = sqrtf(rms / size + eps);
	float scale = 1.0f / rms;
	for (int i = 0; i < size; ++i) o[i] = x[i] * scale * weight[i];
}

/* src/infer.cpp:170-185 */
static void yo_softmax(float *o, const float *x, int size) {
	float score_max = -FLT_MAX;
	for (int i = 0; i < size; ++i)
		if (x[i] > score_max) score_max = x[i];
	float score_sum = 0.0f;
	for (int i = 0; i < size; ++i) {
		o[i] = expf(x[i] - score_max);
		score_sum += o[i];
	}
	for (int i = 0; i < size; ++i) o[i] /= score_sum;
}

static float yo_gelu(float x) { return 0.5f * x * (1.0f + tanhf(0.797885f * (x + 0.044715f * x * x * x))); } /* :187-189 */
static float yo_silu(float x) { return x / (1.0f + expf(-x)); }                                               /* :191-193 */
static float yo_clip(float x, float v) { return x < -v ? -v : (x > v ? v : x); }                             /* :195-197 */

/* src/infer.cpp:200-213: interleaved pairs (i, i+1); frequency index is the position inside the head. */
void yo_rope(float *vec, int d, int head_dim, int pos, float theta, int rotary_dim) {
	for (int i = 0; i < d; i += 2) {
		int j_head = i % head_dim;
		float freq = j_head >= rotary_dim ? 0.f : 1.0f / powf(theta, (float)j_head / (float)rotary_dim);
		float val = pos * freq;
		float fcr = cosf(val), fci = sinf(val);
		float v0 = vec[i], v1 = vec[i + 1];
		vec[i] = v0 * fcr - v1 * fci;
		vec[i + 1] = v0 * fci + v1 * fcr;
	}
}

/* src/infer.cpp:216-248 */
void yo_attn(float *xout, float *atth, const float *qh, const uint16_t *kh, const uint16_t *vh, int head_dim,
             int n_kv_heads, int kv_len) {
	int kv_stride = n_kv_heads * head_dim;
	for (int t = 0; t < kv_len; ++t) {
		float score = 0.0f;
		for (int i = 0; i < head_dim; ++i) score += qh[i] * yo_half_to_float(kh[(size_t)t * kv_stride + i]);
		score /= sqrtf(head_dim);
		atth[t] = score;
	}
	yo_softmax(atth, atth, kv_len);
	for (int i = 0; i < head_dim; ++i) {
		float vi = 0.0f;
		for (int t = 0; t < kv_len; ++t) vi += atth[t] * yo_half_to_float(vh[(size_t)t * kv_stride + i]);
		xout[i] = vi;
	}
}

/* src/infer.cpp:387-403 */
void yo_mha(float *xout, float *att, const uint16_t *kb, const uint16_t *vb, const float *q, int head_dim,
            int kv_len, int max_seq_len, int n_heads, int n_kv_heads) {
	int q_per_kv_head = n_heads / n_kv_heads;
	int h;
#pragma omp parallel for private(h)
	for (h = 0; h < n_heads; h++) {
		int off = (h / q_per_kv_head) * head_dim;
		yo_attn(xout + head_dim * h, att + (size_t)max_seq_len * h, q + head_dim * h, kb + off, vb + off, head_dim,
		        n_kv_heads, kv_len);
	}
}

/* src/infer.cpp:100-132: top-k by repeated arg-max with strict '>' (lowest index wins ties); weights are
 * a softmax over the SELECTED experts only (each exp taken relative to the global max). */
void yo_moe_gate(float *moe_weights, int *active_experts, const float *x, int n_experts, int n_active) {
	float max_val = -FLT_MAX;
	for (int j = 0; j < n_experts; ++j)
		if (x[j] > max_val) max_val = x[j];
	uint64_t mask = 0;
	float wsum = 0.0f;
	for (int k = 0; k < n_active; ++k) {
		int best = -1;
		for (int j = 0; j < n_experts; ++j)
			if ((mask & (1ull << j)) == 0 && (best == -1 || x[j] > x[best])) best = j;
		active_experts[k] = best;
		wsum += expf(x[best] - max_val);
		mask |= 1ull << best;
	}
	for (int k = 0; k < n_active; ++k) moe_weights[k] = expf(x[active_experts[k]] - max_val) / wsum;
}

static size_t yo_wsize(int dtype) { return dtype == YO_F32 ? 4 : dtype == YO_F16 ? 2 : 1; }

/* src/infer.cpp:412-438 */
void yo_ffn(float *xout, const float *x, const void *w1, const void *w2, const void *w3, const float *s1,
            const float *s2, const float *s3, int hidden_dim, int dim, int act, int dtype) {
	float hb[hidden_dim], hb2[hidden_dim];
	yo_matmul(hb, x, w1, s1, dim, hidden_dim, dtype);
	yo_matmul(hb2, x, w3, s3, dim, hidden_dim, dtype);
	for (int i = 0; i < hidden_dim; ++i) hb[i] = (act == YO_GELU ? yo_gelu(hb[i]) : yo_silu(hb[i])) * hb2[i];
	yo_matmul(xout, hb, w2, s2, hidden_dim, dim, dtype);
}

/* src/infer.cpp:254-385 */
void yo_block(yo_model *m, int layer, int pos, int kv_sink, int kv_pos, int kv_len) {
	const yo_config *c = &m->c;
	yo_layer *L = &m->layers[layer];
	int dt = c->weight_dtype;
	int q_dim = c->n_heads * c->head_dim, kv_dim = c->n_kv_heads * c->head_dim;

	yo_rmsnorm(m->xb, m->x, L->rms_att, c->dim, c->norm_eps);
	yo_matmul(m->q, m->xb, L->wq, L->sq, c->dim, q_dim, dt);
	yo_matmul(m->k, m->xb, L->wk, L->sk, c->dim, kv_dim, dt);
	yo_matmul(m->v, m->xb, L->wv, L->sv, c->dim, kv_dim, dt);
	for (int i = 0; i < q_dim; ++i) m->q[i] = yo_clip(m->q[i], c->qkv_clip);
	for (int i = 0; i < kv_dim; ++i) {
		m->k[i] = yo_clip(m->k[i], c->qkv_clip);
		m->v[i] = yo_clip(m->v[i], c->qkv_clip);
	}
	yo_rope(m->q, q_dim, c->head_dim, pos, c->rope_theta, c->rotary_dim);
	yo_rope(m->k, kv_dim, c->head_dim, pos, c->rope_theta, c->rotary_dim);

	uint16_t *kb = L->key_cache, *vb = L->value_cache;
	for (int i = 0; i < kv_dim; ++i) {
		kb[(size_t)kv_pos * kv_dim + i] = yo_float_to_half(m->k[i]);
		vb[(size_t)kv_pos * kv_dim + i] = yo_float_to_half(m->v[i]);
	}
	/* attention sinks: re-rotate each sink key by one position, through fp32 (:303-317) */
	for (int r = 0; r < kv_sink; r++) {
		for (int i = 0; i < kv_dim; ++i) m->k[i] = yo_half_to_float(kb[(size_t)r * kv_dim + i]);
		yo_rope(m->k, kv_dim, c->head_dim, 1, c->rope_theta, c->rotary_dim);
		for (int i = 0; i < kv_dim; i++) kb[(size_t)r * kv_dim + i] = yo_float_to_half(m->k[i]);
	}
	yo_mha(m->xb2, m->att, kb, vb, m->q, c->head_dim, kv_len, c->max_seq_len, c->n_heads, c->n_kv_heads);

	yo_matmul(m->hb, m->xb2, L->wo, L->so, q_dim, c->dim, dt);
	for (int i = 0; i < c->dim; ++i) m->x[i] += m->hb[i];

	yo_rmsnorm(m->xb, m->x, L->rms_ffn, c->dim, c->norm_eps);
	int n_active = 1;
	if (c->n_experts > 0) {
		yo_matmul(m->moe_weights, m->xb, L->moegate, L->sgate, c->dim, c->n_experts, dt);
		yo_moe_gate(m->active_w, m->active_e, m->moe_weights, c->n_experts, c->n_experts_active);
		n_active = c->n_experts_active;
	} else {
		m->active_w[0] = 1.0f;
		m->active_e[0] = 0;
	}
	size_t esz = (size_t)c->dim * c->hidden_dim, ws = yo_wsize(dt);
	for (int k = 0; k < n_active; ++k) {
		int e = m->active_e[k];
		const char *w1 = (const char *)L->w1 + e * esz * ws, *w2 = (const char *)L->w2 + e * esz * ws,
		           *w3 = (const char *)L->w3 + e * esz * ws;
		const float *s1 = L->s1 ? L->s1 + (size_t)e * c->hidden_dim : 0, *s3 = L->s3 ? L->s3 + (size_t)e * c->hidden_dim : 0,
		            *s2 = L->s2 ? L->s2 + (size_t)e * c->dim : 0;
		yo_matmul(m->hb, m->xb, w1, s1, c->dim, c->hidden_dim, dt);
		yo_matmul(m->hb2, m->xb, w3, s3, c->dim, c->hidden_dim, dt);
		for (int i = 0; i < c->hidden_dim; ++i)
			m->hb[i] = (c->act == YO_GELU ? yo_gelu(m->hb[i]) : yo_silu(m->hb[i])) * m->hb2[i];
		yo_matmul(m->xb2, m->hb, w2, s2, c->hidden_dim, c->dim, dt);
		float ew = m->active_w[k];
		for (int i = 0; i < c->dim; ++i) m->x[i] += m->xb2[i] * ew;
	}
}

/* src/infer.cpp:443-524; KV ring buffer with KV_SINKS = 2 attention sinks (src/model.h:12, :483-485). */
void yo_forward(yo_model *m, int token, int pos, int mode) {
	const yo_config *c = &m->c;
	for (int i = 0; i < c->dim; ++i) {
		size_t o = (size_t)token * c->dim + i;
		if (c->weight_dtype == YO_F32) m->x[i] = ((const float *)m->embed)[o];
		else if (c->weight_dtype == YO_F16) m->x[i] = yo_half_to_float(((const uint16_t *)m->embed)[o]);
		else m->x[i] = yo_e4m3_to_float(((const uint8_t *)m->embed)[o]) * (m->embed_scale ? m->embed_scale[token] : 1.0f);
	}
	int kv_sink = pos >= c->max_seq_len ? 2 : 0;
	int kv_pos = kv_sink + (pos - kv_sink) % (c->max_seq_len - kv_sink);
	int kv_len = pos >= c->max_seq_len ? c->max_seq_len : pos + 1;
	for (int l = 0; l < c->n_layers; ++l) yo_block(m, l, pos, kv_sink, kv_pos, kv_len);
	if (mode == 0) return;
	yo_rmsnorm(m->x, m->x, m->rms_final, c->dim, c->norm_eps);
	yo_matmul(m->logits, m->x, m->wcls, m->wcls_scale, c->dim, c->vocab_size, c->weight_dtype);
}

/* src/sampler.cpp:27-38: first index of the maximum */
int yo_argmax(const float *logits, int n) {
	int argmax = 0;
	float max_val = -FLT_MAX;
	for (int i = 0; i < n; ++i)
		if (logits[i] > max_val) {
			max_val = logits[i];
			argmax = i;
		}
	return argmax;
}
