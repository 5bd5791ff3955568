// C ABI layer (3): host-pointer test entry points (reference semantics: src/infer.cu:890-1007) and
// opaque handles over Model / InferenceState for ctypes / cgo / JNI style bindings.
#include "yalm_api.h"
#include "tokenizer.h"

#include <cuda_runtime.h>

#include <algorithm>
#include <cstdio>
#include <cstring>
#include <memory>
#include <string>
#include <vector>

extern "C" int yalm_attn_generic(float *out, float *att, const uint16_t *key_cache, const uint16_t *value_cache,
                                 const float *q, int n_heads, int n_kv_heads, int head_dim, int max_seq_len,
                                 int kv_len, const int *ctl, yalm_stream_t s);
extern "C" int yalm_attn_is_fast_shape(int n_heads, int n_kv_heads, int head_dim);
extern "C" int yalm_stream_probe(size_t bytes, int iters, double *usec);

namespace {

int have_gpu() {
	int n = 0;
	return cudaGetDeviceCount(&n) == cudaSuccess && n > 0;
}

// RAII device buffer for the test entry points (the reference leaks these, "all cuda uploads leak forever").
struct DevBuf {
	void *p = nullptr;
	DevBuf() = default;
	DevBuf(const void *host, size_t bytes) { alloc(bytes, host); }
	int alloc(size_t bytes, const void *host = nullptr) {
		cudaError_t e = cudaMalloc(&p, bytes ? bytes : 16);
		if (e != cudaSuccess) return (int)e;
		if (host) e = cudaMemcpy(p, host, bytes, cudaMemcpyHostToDevice);
		else e = cudaMemset(p, 0, bytes ? bytes : 16);
		return (int)e;
	}
	~DevBuf() { cudaFree(p); }
	template <typename T> T *as() const { return static_cast<T *>(p); }
};

size_t elem_bytes(int dtype) {
	return dtype == YALM_F32 ? 4 : dtype == YALM_F16 ? 2 : 1;
}

#define RET_IF(x)                                                                                             \
	do {                                                                                                      \
		int rc__ = (int)(x);                                                                                  \
		if (rc__ != 0) return rc__;                                                                           \
	} while (0)

} // namespace

// ---- (3a) test entry points --------------------------------------------------------------------------
extern "C" int yalm_matmul_host(float *xout, const float *x, const void *w, const float *scale, int n, int d,
                                int dtype) {
	if (!have_gpu()) return YALM_ENOGPU;
	if (!xout || !x || !w || n <= 0 || d <= 0) return YALM_EINVAL;
	DevBuf dx(x, sizeof(float) * n), dw(w, elem_bytes(dtype) * (size_t)n * d), dout, ds;
	RET_IF(dout.alloc(sizeof(float) * d));
	if (scale) RET_IF(ds.alloc(sizeof(float) * d, scale));
	RET_IF(yalm_gemv(dout.as<float>(), dx.as<float>(), dw.p, scale ? ds.as<float>() : nullptr, n, d, dtype, nullptr));
	RET_IF(cudaDeviceSynchronize());
	return (int)cudaMemcpy(xout, dout.p, sizeof(float) * d, cudaMemcpyDeviceToHost);
}

extern "C" int yalm_mha_host(float *xout, float *att, const uint16_t *kb, const uint16_t *vb, const float *q,
                             int head_dim, int kv_len, int max_seq_len, int n_heads, int n_kv_heads) {
	if (!have_gpu()) return YALM_ENOGPU;
	if (!xout || !kb || !vb || !q || kv_len < 1 || kv_len > max_seq_len || n_kv_heads < 1 || n_heads % n_kv_heads)
		return YALM_EINVAL;
	const size_t kv_elems = (size_t)max_seq_len * n_kv_heads * head_dim, kv_bytes = sizeof(uint16_t) * kv_elems;
	// reference layout (t, kv head, d) -> the device's head-major ring (kv head, t, d)
	std::vector<uint16_t> kh(kv_elems), vh(kv_elems);
	for (int t = 0; t < max_seq_len; ++t)
		for (int h = 0; h < n_kv_heads; ++h) {
			const size_t src = ((size_t)t * n_kv_heads + h) * head_dim, dst = ((size_t)h * max_seq_len + t) * head_dim;
			memcpy(&kh[dst], kb + src, sizeof(uint16_t) * head_dim);
			memcpy(&vh[dst], vb + src, sizeof(uint16_t) * head_dim);
		}
	DevBuf dk(kh.data(), kv_bytes), dv(vh.data(), kv_bytes), dq(q, sizeof(float) * n_heads * head_dim), dout, datt, dctl, dws;
	RET_IF(dout.alloc(sizeof(float) * n_heads * head_dim));
	int ctl[YALM_CTL_INTS] = {0, kv_len - 1, 0, kv_len - 1, kv_len, 0, 0, 0};
	RET_IF(dctl.alloc(sizeof(ctl), ctl));
	const bool fast = yalm_attn_is_fast_shape(n_heads, n_kv_heads, head_dim);
	if (att || !fast) {
		// materialised, normalised attention rows, as the reference returns them in `att`
		RET_IF(datt.alloc(sizeof(float) * (size_t)n_heads * max_seq_len));
		RET_IF(yalm_attn_generic(dout.as<float>(), datt.as<float>(), dk.as<uint16_t>(), dv.as<uint16_t>(), dq.as<float>(),
		                         n_heads, n_kv_heads, head_dim, max_seq_len, kv_len, nullptr, nullptr));
		RET_IF(cudaDeviceSynchronize());
		if (att) RET_IF(cudaMemcpy(att, datt.p, sizeof(float) * (size_t)n_heads * max_seq_len, cudaMemcpyDeviceToHost));
	}
	if (fast) { // xout always comes from the production flash-decode kernel when the shape allows
		RET_IF(dws.alloc(yalm_attn_workspace_bytes(n_heads, n_kv_heads, head_dim)));
		RET_IF(yalm_attn_decode(dout.as<float>(), dk.as<uint16_t>(), dv.as<uint16_t>(), dq.as<float>(), dws.p, n_heads,
		                        n_kv_heads, head_dim, max_seq_len, dctl.as<int>(), nullptr));
		RET_IF(cudaDeviceSynchronize());
	}
	return (int)cudaMemcpy(xout, dout.p, sizeof(float) * n_heads * head_dim, cudaMemcpyDeviceToHost);
}

extern "C" int yalm_ffn_host(float *xout, const float *x, const void *w1, const void *w2, const void *w3,
                             const float *s1, const float *s2, const float *s3, int hidden_dim, int dim, int act,
                             int dtype) {
	if (!have_gpu()) return YALM_ENOGPU;
	if (!xout || !x || !w1 || !w2 || !w3 || hidden_dim <= 0 || dim <= 0) return YALM_EINVAL;
	const size_t wb = elem_bytes(dtype) * (size_t)hidden_dim * dim;
	DevBuf dx(x, sizeof(float) * dim), d1(w1, wb), d2(w2, wb), d3(w3, wb), dhb, dout, ds1, ds2, ds3;
	RET_IF(dhb.alloc(sizeof(float) * hidden_dim));
	RET_IF(dout.alloc(sizeof(float) * dim));
	if (s1 && s2 && s3) {
		RET_IF(ds1.alloc(sizeof(float) * hidden_dim, s1));
		RET_IF(ds3.alloc(sizeof(float) * hidden_dim, s3));
		RET_IF(ds2.alloc(sizeof(float) * dim, s2));
	}
	// ffn_cuda has no norm in front of it (src/infer.cu:959-993): rms_weight = NULL selects the
	// un-normalised variant of the fused w1/w3/GLU kernel; then the w2 GEMV with no residual.
	RET_IF(yalm_ffn_glu(dhb.as<float>(), dx.as<float>(), nullptr, d1.p, d3.p, ds1.as<float>(), ds3.as<float>(), nullptr,
	                    nullptr, dim, hidden_dim, act, 0.f, dtype, 0, 0, nullptr, nullptr, nullptr, nullptr));
	RET_IF(yalm_gemv_residual(dout.as<float>(), nullptr, dhb.as<float>(), d2.p, ds2.as<float>(), hidden_dim, dim, dtype, 1,
	                          nullptr, nullptr, nullptr));
	RET_IF(cudaDeviceSynchronize());
	return (int)cudaMemcpy(xout, dout.p, sizeof(float) * dim, cudaMemcpyDeviceToHost);
}

// C++ names of the reference's test entry points (src/model.h:353-384).
template <typename T> struct DtypeOf;
template <> struct DtypeOf<float> { static constexpr int v = YALM_F32; };
template <> struct DtypeOf<f16_t> { static constexpr int v = YALM_F16; };
template <> struct DtypeOf<uint8_t> { static constexpr int v = YALM_F8E4M3; };

static void die_on(int rc, const char *what) {
	if (rc != 0) {
		fprintf(stderr, "yalm_b200: %s failed: %s (%d)\n", what, yalm_error_string(rc), rc);
		abort();
	}
}

void mha_cuda(float *xout, float *att, f16_t *kb, f16_t *vb, float *q, int head_dim, int kv_len, int max_seq_len,
              int n_heads, int n_kv_heads) {
	die_on(yalm_mha_host(xout, att, kb, vb, q, head_dim, kv_len, max_seq_len, n_heads, n_kv_heads), "mha_cuda");
}
template <typename T> void matmul_cuda(float *xout, float *x, T *w, int n, int d) {
	die_on(yalm_matmul_host(xout, x, w, nullptr, n, d, DtypeOf<T>::v), "matmul_cuda");
}
template <typename T>
void ffn_cuda(float *xout, float *x, T *w1, T *w2, T *w3, int hidden_dim, int dim, ActivationType act) {
	die_on(yalm_ffn_host(xout, x, w1, w2, w3, nullptr, nullptr, nullptr, hidden_dim, dim,
	                     act == ActivationType::SILU ? YALM_ACT_SILU : YALM_ACT_GELU, DtypeOf<T>::v),
	       "ffn_cuda");
}
template void matmul_cuda<float>(float *, float *, float *, int, int);
template void matmul_cuda<f16_t>(float *, float *, f16_t *, int, int);
template void matmul_cuda<uint8_t>(float *, float *, uint8_t *, int, int);
template void ffn_cuda<float>(float *, float *, float *, float *, float *, int, int, ActivationType);
template void ffn_cuda<f16_t>(float *, float *, f16_t *, f16_t *, f16_t *, int, int, ActivationType);
template void ffn_cuda<uint8_t>(float *, float *, uint8_t *, uint8_t *, uint8_t *, int, int, ActivationType);

// ---- kernel bench: what `./build/test -k <name>` prints (reference shapes, src/test.cpp:307-359) --------
extern "C" int yalm_bench_op(const char *name, int iters, int warmup, double *usec, double *bytes) {
	if (!have_gpu()) return YALM_ENOGPU;
	if (!name || iters < 1 || !usec) return YALM_EINVAL;
	const std::string k(name);
	if (k == "stream") { // ceiling of the megakernel's TMA ring: pure streaming read of 4 GiB
		const double b = 4.0 * (1u << 30);
		if (bytes) *bytes = b;
		return yalm_stream_probe((size_t)b, iters, usec);
	}
	const int head_dim = 128, n_heads = 32, dim = head_dim * n_heads, n_kv_heads = 8, kv_len = 4096;
	int hidden = 14336;
	cudaStream_t st;
	RET_IF(cudaStreamCreate(&st));
	cudaEvent_t e0, e1;
	cudaEventCreate(&e0);
	cudaEventCreate(&e1);
	// a buffer larger than L2 to flush it between iterations
	DevBuf flush;
	const size_t flush_bytes = 256u << 20;
	RET_IF(flush.alloc(flush_bytes));
	double algo_bytes = 0;
	float total_ms = 0.f;
	auto timed = [&](auto &&fn) -> int {
		for (int i = 0; i < warmup + iters; ++i) {
			RET_IF(cudaMemsetAsync(flush.p, i, flush_bytes, st));
			RET_IF(cudaEventRecord(e0, st));
			RET_IF(fn());
			RET_IF(cudaEventRecord(e1, st));
			RET_IF(cudaStreamSynchronize(st));
			float ms = 0.f;
			cudaEventElapsedTime(&ms, e0, e1);
			if (i >= warmup) total_ms += ms;
		}
		return 0;
	};
	int rc = 0;
	if (k == "matmul" || k == "matmul-wide") {
		if (k == "matmul-wide") hidden = 32000;
		DevBuf w, x, out;
		RET_IF(w.alloc(2 * (size_t)dim * hidden));
		RET_IF(cudaMemset(w.p, 0x3c, 2 * (size_t)dim * hidden)); // finite fp16 pattern
		RET_IF(x.alloc(4 * (size_t)dim));
		RET_IF(out.alloc(4 * (size_t)hidden));
		algo_bytes = 2.0 * dim * hidden;
		rc = timed([&] { return yalm_gemv(out.as<float>(), x.as<float>(), w.p, nullptr, dim, hidden, YALM_F16, st); });
	} else if (k == "mha") {
		const size_t kvb = 2 * (size_t)kv_len * n_kv_heads * head_dim;
		DevBuf kb, vb, q, out, ws, ctl;
		RET_IF(kb.alloc(kvb));
		RET_IF(vb.alloc(kvb));
		RET_IF(cudaMemset(kb.p, 0x2c, kvb));
		RET_IF(cudaMemset(vb.p, 0x2c, kvb));
		RET_IF(q.alloc(4 * (size_t)dim));
		RET_IF(out.alloc(4 * (size_t)dim));
		RET_IF(ws.alloc(yalm_attn_workspace_bytes(n_heads, n_kv_heads, head_dim)));
		int h[YALM_CTL_INTS] = {0, kv_len - 1, 0, kv_len - 1, kv_len, 0, 0, 0};
		RET_IF(ctl.alloc(sizeof(h), h));
		algo_bytes = 2.0 * kvb;
		rc = timed([&] {
			return yalm_attn_decode(out.as<float>(), kb.as<uint16_t>(), vb.as<uint16_t>(), q.as<float>(), ws.p, n_heads,
			                        n_kv_heads, head_dim, kv_len, ctl.as<int>(), st);
		});
	} else if (k == "ffn") {
		const size_t wb = 2 * (size_t)dim * hidden;
		DevBuf w1, w2, w3, x, hb, out;
		RET_IF(w1.alloc(wb));
		RET_IF(w2.alloc(wb));
		RET_IF(w3.alloc(wb));
		RET_IF(cudaMemset(w1.p, 0x1c, wb));
		RET_IF(cudaMemset(w2.p, 0x1c, wb));
		RET_IF(cudaMemset(w3.p, 0x1c, wb));
		RET_IF(x.alloc(4 * (size_t)dim));
		RET_IF(hb.alloc(4 * (size_t)hidden));
		RET_IF(out.alloc(4 * (size_t)dim));
		algo_bytes = 3.0 * wb;
		rc = timed([&] {
			int r = yalm_ffn_glu(hb.as<float>(), x.as<float>(), nullptr, w1.p, w3.p, nullptr, nullptr, nullptr, nullptr,
			                     dim, hidden, YALM_ACT_GELU, 0.f, YALM_F16, 0, 0, nullptr, nullptr, nullptr, st);
			if (r) return r;
			return yalm_gemv_residual(out.as<float>(), nullptr, hb.as<float>(), w2.p, nullptr, hidden, dim, YALM_F16, 1,
			                          nullptr, nullptr, st);
		});
	} else {
		rc = YALM_EINVAL;
	}
	cudaEventDestroy(e0);
	cudaEventDestroy(e1);
	cudaStreamDestroy(st);
	if (rc) return rc;
	*usec = 1000.0 * total_ms / iters;
	if (bytes) *bytes = algo_bytes;
	return 0;
}

// ---- (3b) opaque handles --------------------------------------------------------------------------------
struct yalm_model {
	YALMData data;
	std::unique_ptr<Model> model;
};
struct yalm_state {
	std::unique_ptr<InferenceState> state;
};

static int open_common(yalm_model *h, int context, yalm_model **out) {
	for (const char *k : {"dim", "hidden_dim", "head_dim", "n_layers", "n_heads", "n_kv_heads", "vocab_size",
	                      "max_seq_len", "rope_theta", "rotary_dim", "dtype"})
		if (!h->data.metadata.contains(k)) {
			fprintf(stderr, "yalm_b200: metadata key %s missing\n", k);
			delete h;
			return YALM_EFORMAT;
		}
	h->model = std::make_unique<Model>(h->data, context);
	*out = h;
	return 0;
}

extern "C" int yalm_model_open_file(const char *path, int context, yalm_model **out) {
	if (!path || !out) return YALM_EINVAL;
	auto *h = new yalm_model();
	if (h->data.from_file(path) != 0) {
		delete h;
		return YALM_EFORMAT;
	}
	return open_common(h, context, out);
}

extern "C" int yalm_model_open_memory(const void *image, size_t size, int context, yalm_model **out) {
	if (!image || !out) return YALM_EINVAL;
	auto *h = new yalm_model();
	if (h->data.from_memory(image, size) != 0) {
		delete h;
		return YALM_EFORMAT;
	}
	return open_common(h, context, out);
}

extern "C" int yalm_model_cuda(yalm_model *m, int device, int tp_rank, int tp_size) {
	if (!m) return YALM_EINVAL;
	if (!have_gpu()) return YALM_ENOGPU;
	m->model->cuda(device, tp_rank, tp_size);
	return 0;
}

extern "C" void yalm_model_close(yalm_model *m) {
	delete m;
}

extern "C" void yalm_model_config(const yalm_model *m, int *out) {
	const Config &c = *m->model->config;
	int v[16] = {c.dim,        c.hidden_dim,  c.head_dim,    c.n_layers,   c.n_heads,          c.n_kv_heads,
	             c.vocab_size, c.max_seq_len, c.rotary_dim,  c.n_experts,  c.n_experts_active, (int)c.weight_dtype,
	             c.act == ActivationType::SILU ? 1 : 0, c.tp_rank, c.tp_size, 0};
	memcpy(out, v, sizeof(v));
}

extern "C" size_t yalm_model_device_bytes(const yalm_model *m) {
	return m->model->device_bytes();
}

extern "C" void yalm_model_load_info(const yalm_model *m, int *out2) {
	out2[0] = m->model->device_allocations();
	out2[1] = m->model->upload_was_pinned() ? 1 : 0;
}

extern "C" int yalm_state_new(yalm_model *m, yalm_state **out) {
	if (!m || !out) return YALM_EINVAL;
	if (!have_gpu() || m->model->device() != Device::CUDA) return YALM_ENOGPU;
	auto *s = new yalm_state();
	s->state = std::make_unique<InferenceState>(m->model->config);
	s->state->cuda();
	*out = s;
	return 0;
}

extern "C" void yalm_state_free(yalm_state *s) {
	delete s;
}

extern "C" int yalm_forward(yalm_model *m, yalm_state *s, int token, int pos, int mode) {
	if (!m || !s) return YALM_EINVAL;
	if (m->model->device() != Device::CUDA) return YALM_ENOGPU;
	if (token < 0 || token >= m->model->config->vocab_size || pos < 0) return YALM_EINVAL;
	m->model->forward(*s->state, token, pos,
	                  mode == YALM_MODE_HYDRATE ? InferenceMode::HYDRATE_KV_CACHE : InferenceMode::OUTPUT_LOGITS);
	return 0;
}

extern "C" float *yalm_logits(yalm_state *s) {
	return s->state->logits();
}

extern "C" int yalm_generate_greedy(yalm_model *m, yalm_state *s, int token, int pos, int n_gen, int *tokens_out,
                                    float *logits_out, float *device_ms) {
	if (!m || !s) return YALM_EINVAL;
	return m->model->generate_greedy(*s->state, token, pos, n_gen, tokens_out, logits_out, device_ms);
}

extern "C" int yalm_model_batch_capacity(const yalm_model *m) {
	return m ? m->model->batch_capacity() : 0;
}

extern "C" int yalm_forward_batch(yalm_model *m, yalm_state *s, const int *tokens, int n, int pos0, const int *targets, float *probs_out) {
	if (!m || !s) return YALM_EINVAL;
	return m->model->forward_batch(*s->state, tokens, n, pos0, targets, probs_out);
}

extern "C" int yalm_generate(yalm_model *m, yalm_state *s, int token, int pos, int n_gen, float temperature,
                             unsigned long long seed, int *tokens_out, float *logits_out, float *device_ms) {
	if (!m || !s) return YALM_EINVAL;
	return m->model->generate(*s->state, token, pos, n_gen, temperature, seed, tokens_out, logits_out, device_ms);
}

extern "C" void yalm_state_debug_set_launch_seq(yalm_state *s, unsigned seq) {
	if (s) s->state->set_mega_seq(seq);
}

extern "C" int yalm_state_read(yalm_model *m, yalm_state *sh, int which, int layer, void *dst, size_t bytes) {
	if (!m || !sh || !dst) return YALM_EINVAL;
	InferenceState &s = *sh->state;
	const void *src = nullptr;
	switch (which) {
	case 0: src = s.x(); break;
	case 1: src = s.xb2(); break;
	case 2: src = s.hb(); break;
	case 3: src = s.q(); break;
	case 4:
	case 5:
		if (layer < 0 || layer >= (int)m->model->blocks.size()) return YALM_EINVAL;
		src = which == 4 ? (const void *)m->model->blocks[layer]->key_cache()
		                 : (const void *)m->model->blocks[layer]->value_cache();
		break;
	case 6: src = s.active_experts(); break;
	case 7: src = s.active_experts_weights(); break;
	default: return YALM_EINVAL;
	}
	RET_IF(cudaStreamSynchronize(s.stream()));
	if (which == 4 || which == 5) {
		// device ring is head-major; the tap returns the reference layout (t, kv head, d)
		const Config &c = *m->model->config;
		const int H = c.local_kv_heads(), D = c.head_dim, T = c.max_seq_len;
		const size_t n = (size_t)H * T * D;
		if (bytes > n * sizeof(uint16_t) || bytes % (sizeof(uint16_t) * H * D)) return YALM_EINVAL;
		std::vector<uint16_t> tmp(n);
		RET_IF(cudaMemcpy(tmp.data(), src, n * sizeof(uint16_t), cudaMemcpyDeviceToHost));
		const int rows = (int)(bytes / (sizeof(uint16_t) * H * D));
		uint16_t *out = static_cast<uint16_t *>(dst);
		for (int t = 0; t < rows; ++t)
			for (int h = 0; h < H; ++h)
				memcpy(out + ((size_t)t * H + h) * D, &tmp[((size_t)h * T + t) * D], sizeof(uint16_t) * D);
		return 0;
	}
	return (int)cudaMemcpy(dst, src, bytes, cudaMemcpyDeviceToHost);
}

extern "C" int yalm_forward_launches(const yalm_model *m, int mode) {
	return m->model->launches_per_forward(mode == YALM_MODE_HYDRATE ? InferenceMode::HYDRATE_KV_CACHE
	                                                                : InferenceMode::OUTPUT_LOGITS);
}

extern "C" int yalm_model_uses_megakernel(const yalm_model *m) {
	return m && m->model->uses_megakernel() ? 1 : 0;
}

extern "C" int yalm_model_mega_info(const yalm_model *m, int *out8) {
	if (!m || !out8) return YALM_EINVAL;
	return m->model->mega_info(out8);
}

extern "C" int yalm_mega_profile(yalm_model *m, yalm_state *s, int token, int pos, long long *out, size_t count, int *grid) {
	if (!m || !s) return YALM_EINVAL;
	return m->model->mega_profile(*s->state, token, pos, out, count, grid);
}

extern "C" int yalm_profile_ffn(yalm_model *m, yalm_state *s, int token, int pos, int n_steps, double *mean_us) {
	if (!m || !s) return YALM_EINVAL;
	return m->model->profile_ffn(*s->state, token, pos, n_steps, mean_us);
}

// ---- tokenizer taps (host only; the CLI's Tokenizer over an in-memory .yalm image) -------------------------------------
extern "C" int yalm_tokenize(const void *image, size_t size, const char *text, int add_bos, int *out, int cap) {
	if (!image || !text || !out || cap < 0) return YALM_EINVAL;
	YALMData d;
	if (d.from_memory(image, size) != 0 || !d.metadata.contains("bos_token_id") || !d.metadata.contains("eos_token_id")) return YALM_EFORMAT;
	Tokenizer t(d);
	const std::vector<int> ids = t.encode(text, add_bos != 0);
	for (size_t i = 0; i < ids.size() && (int)i < cap; ++i) out[i] = ids[i];
	return (int)ids.size();
}

extern "C" int yalm_detokenize(const void *image, size_t size, const int *ids, int n, char *out, int cap) {
	if (!image || !ids || !out || n < 0 || cap < 1) return YALM_EINVAL;
	YALMData d;
	if (d.from_memory(image, size) != 0 || !d.metadata.contains("bos_token_id") || !d.metadata.contains("eos_token_id")) return YALM_EFORMAT;
	Tokenizer t(d);
	std::string s;
	for (int i = 0; i < n; ++i) {
		if (ids[i] < 0 || ids[i] >= (int)t.vocab.size()) return YALM_EINVAL;
		s += t.decode_one(i > 0 ? ids[i - 1] : -1, ids[i]);
	}
	const int len = (int)s.size();
	memcpy(out, s.data(), std::min(len, cap - 1));
	out[std::min(len, cap - 1)] = 0;
	return len;
}
