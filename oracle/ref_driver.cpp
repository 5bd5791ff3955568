// TEST INFRASTRUCTURE — NOT PRODUCT CODE.
//
// C-ABI harness around the UNMODIFIED reference (xiaguan/yalm) objects, linked
// into oracle/_ref/libyalm_ref.so by oracle/Makefile.  Only tests/,
// __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs
// may load it.  The reference's `main` prints decoded text only
// (src/main.cpp:102-112), so parity checks need this harness to reach
// Model::forward (src/model.cpp:396-407) and InferenceState::logits()
// (src/model.h:139-141) directly.
//
// Built with -fno-access-control so that the dense-model crash of the
// reference CPU path can be worked around WITHOUT patching reference sources:
// _block_cpu stores through active_experts()/active_experts_weights() when
// n_experts == 0 (src/infer.cpp:350-353) but the buffers are only allocated
// for n_experts > 0 (src/model.cpp:279-283).  ref_open() allocates the two
// one-element buffers after constructing the state.

#include "model.h"
#include "sampler.h"
#include "tokenizer.h"

#include <cstdio>
#include <cstring>
#include <memory>
#include <omp.h>
#include <string>
#include <vector>

namespace {
struct RefHandle {
	YALMData data;
	std::unique_ptr<Model> model;
	std::unique_ptr<InferenceState> state;
	std::unique_ptr<Sampler> sampler;
};
} // namespace

extern "C" {

// Opens a .yalm file with the reference loader and builds Model +
// InferenceState on the CPU.  `context` is the reference's -T value
// (0 = min(metadata max_seq_len, 4096), src/model.cpp:31-36).
void *ref_open(const char *path, int context) {
	auto *h = new RefHandle();
	if (h->data.from_file(path) != 0) {
		delete h;
		return nullptr;
	}
	h->model = std::make_unique<Model>(h->data, context);
	h->state = std::make_unique<InferenceState>(h->model->config);
	if (h->model->config->n_experts == 0) {
		h->state->_active_experts = new int[1]();
		h->state->_active_experts_weights = new float[1]();
	}
	h->sampler = std::make_unique<Sampler>(h->model->config, 0);
	return h;
}

void ref_close(void *hv) {
	delete static_cast<RefHandle *>(hv);
}

// out[0..15]: dim, hidden_dim, head_dim, n_layers, n_heads, n_kv_heads,
// vocab_size, max_seq_len, rotary_dim, n_experts, n_experts_active,
// weight dtype enum (src/codec.h:15-25), act (0 gelu, 1 silu).
void ref_config(void *hv, int *out) {
	const Config &c = *static_cast<RefHandle *>(hv)->model->config;
	int v[] = {c.dim,        c.hidden_dim,  c.head_dim,   c.n_layers,  c.n_heads,
	           c.n_kv_heads, c.vocab_size,  c.max_seq_len, c.rotary_dim, c.n_experts,
	           c.n_experts_active, (int)c.weight_dtype, (int)c.act};
	memcpy(out, v, sizeof(v));
}

// Moves model + state to the GPU through the reference's own .cuda() calls
// (src/model.cpp:323-345, 380-394).  GPU box only.
void ref_cuda(void *hv) {
	auto *h = static_cast<RefHandle *>(hv);
	h->model->cuda();
	h->state->cuda();
}

// mode: 0 = HYDRATE_KV_CACHE, 1 = OUTPUT_LOGITS (src/model.h:28-31).
void ref_forward(void *hv, int token, int pos, int mode) {
	auto *h = static_cast<RefHandle *>(hv);
	h->model->forward(*h->state, token, pos,
	                  mode ? InferenceMode::OUTPUT_LOGITS : InferenceMode::HYDRATE_KV_CACHE);
}

float *ref_logits(void *hv) {
	return static_cast<RefHandle *>(hv)->state->logits();
}

// Sampler::sample_argmax (src/sampler.cpp:27-38): first index of the maximum.
int ref_argmax(void *hv) {
	auto *h = static_cast<RefHandle *>(hv);
	return h->sampler->sample_argmax(*h->state);
}

float ref_sample_prob(void *hv, int index) {
	auto *h = static_cast<RefHandle *>(hv);
	return h->sampler->sample_prob(index, *h->state);
}

// Host-side activation buffers of the CPU state (valid only before ref_cuda):
// which: 0 x, 1 xb, 2 xb2, 3 hb, 4 q, 5 k, 6 v, 7 att, 8 moe_weights,
// 9 active_experts_weights.  10 returns active_experts (int*).
void *ref_state_buffer(void *hv, int which) {
	InferenceState &s = *static_cast<RefHandle *>(hv)->state;
	switch (which) {
	case 0: return s.x();
	case 1: return s.xb();
	case 2: return s.xb2();
	case 3: return s.hb();
	case 4: return s.q();
	case 5: return s.k();
	case 6: return s.v();
	case 7: return s.att();
	case 8: return s.moe_weights();
	case 9: return s.active_experts_weights();
	case 10: return s.active_experts();
	}
	return nullptr;
}

// KV cache of one layer (host pointers, CPU state only). which: 0 key, 1 value.
void *ref_kv_cache(void *hv, int layer, int which) {
	auto *h = static_cast<RefHandle *>(hv);
	auto &b = h->model->blocks[layer];
	return which == 0 ? (void *)b->key_cache() : (void *)b->value_cache();
}

// Greedy loop exactly as run_completion does it (src/main.cpp:91-112) minus
// tokenizer/printing: hydrate on prompt[0..n_prompt-2], OUTPUT_LOGITS on the
// last prompt token, then n_gen × (argmax, forward).  Writes the generated
// ids to out_tokens and, if logits_out != NULL, the logits that produced each
// generated token (n_gen × vocab fp32).  Returns decode-only seconds.
double ref_generate(void *hv, const int *prompt, int n_prompt, int n_gen, int *out_tokens, float *logits_out) {
	auto *h = static_cast<RefHandle *>(hv);
	const int vocab = h->model->config->vocab_size;
	for (int pos = 0; pos < n_prompt; pos++) {
		h->model->forward(*h->state, prompt[pos], pos,
		                  pos + 1 == n_prompt ? InferenceMode::OUTPUT_LOGITS : InferenceMode::HYDRATE_KV_CACHE);
	}
	double t0 = omp_get_wtime();
	for (int i = 0; i < n_gen; i++) {
		int tok = h->sampler->sample_argmax(*h->state);
		out_tokens[i] = tok;
		if (logits_out) {
			memcpy(logits_out + (size_t)i * vocab, h->state->logits(), sizeof(float) * vocab);
		}
		if (i + 1 < n_gen) {
			h->model->forward(*h->state, tok, n_prompt + i, InferenceMode::OUTPUT_LOGITS);
		}
	}
	return omp_get_wtime() - t0;
}

// The reference's "exposed for tests" free functions (src/model.h:353-384).
void ref_attn(float *xout, float *atth, float *qh, uint16_t *kh, uint16_t *vh, int head_dim, int n_kv_heads,
              int kv_len) {
	attn(xout, atth, qh, kh, vh, head_dim, n_kv_heads, kv_len);
}
void ref_mha_cpu(float *xout, float *att, uint16_t *kb, uint16_t *vb, float *q, int head_dim, int kv_len,
                 int max_seq_len, int n_heads, int n_kv_heads) {
	mha_cpu(xout, att, kb, vb, q, head_dim, kv_len, max_seq_len, n_heads, n_kv_heads);
}
void ref_matmul_cpu_f32(float *xout, float *x, float *w, int n, int d) {
	matmul_cpu(xout, x, w, n, d);
}
void ref_matmul_cpu_f16(float *xout, float *x, uint16_t *w, int n, int d) {
	matmul_cpu(xout, x, w, n, d);
}
void ref_ffn_cpu(float *xout, float *x, float *w1, float *w2, float *w3, int hidden_dim, int dim, int act) {
	ffn_cpu(xout, x, w1, w2, w3, hidden_dim, dim, act ? ActivationType::SILU : ActivationType::GELU);
}
// Reference CUDA entry points (GPU box only) — "the kernel to beat".
void ref_matmul_cuda_f32(float *xout, float *x, float *w, int n, int d) {
	matmul_cuda<float>(xout, x, w, n, d);
}
void ref_matmul_cuda_f16(float *xout, float *x, uint16_t *w, int n, int d) {
	matmul_cuda<f16_t>(xout, x, w, n, d);
}
void ref_mha_cuda(float *xout, float *att, uint16_t *kb, uint16_t *vb, float *q, int head_dim, int kv_len,
                  int max_seq_len, int n_heads, int n_kv_heads) {
	mha_cuda(xout, att, kb, vb, q, head_dim, kv_len, max_seq_len, n_heads, n_kv_heads);
}

int ref_omp_threads() {
	return omp_get_max_threads();
}

// torchrun exports OMP_NUM_THREADS=1 to its workers; the reference arm of bench.py sets the team size explicitly.
void ref_set_threads(int n) {
	if (n > 0) omp_set_num_threads(n);
}

// The reference Tokenizer (src/tokenizer.cpp:57-94, 44-55) on the opened checkpoint: ids of `text`, and the
// concatenated decode_one() pieces of `ids`.  Return the count / byte length.
int ref_tokenize(void *hv, const char *text, int add_bos, int *out, int cap) {
	auto *h = static_cast<RefHandle *>(hv);
	Tokenizer t(h->data);
	std::vector<int> ids = t.encode(text, add_bos != 0);
	for (size_t i = 0; i < ids.size() && (int)i < cap; i++) out[i] = ids[i];
	return (int)ids.size();
}
int ref_detokenize(void *hv, const int *ids, int n, char *out, int cap) {
	auto *h = static_cast<RefHandle *>(hv);
	Tokenizer t(h->data);
	std::string s;
	for (int i = 0; i < n; i++) s += t.decode_one(i > 0 ? ids[i - 1] : -1, ids[i]);
	int len = (int)s.size() < cap - 1 ? (int)s.size() : cap - 1;
	memcpy(out, s.data(), len);
	out[len] = 0;
	return (int)s.size();
}

} // extern "C"
