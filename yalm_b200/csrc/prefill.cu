// Batched prefill / perplexity: n consecutive positions through the whole model in one pass.
// The reference walks a prompt token by token (src/main.cpp:91-97, 174-184); here the projections of a chunk are dense
// GEMMs on the tcgen05 tensor cores (prefill_gemm.cu), attention is a causal flash-attention over the head-major KV ring
// (tensor-core mma.sync m16n8k16, online softmax in fp32), and the KV rows of the whole chunk are appended at once.
// Arithmetic definitions are the decode path's (rmsnorm association order, interleaved-pair RoPE with accurate
// powf/cosf/sinf, clip, gelu/silu, fp16 KV rows); what differs is the summation order inside the GEMMs and that the GEMM
// inputs (normalised x, q, attention output, GLU output) are rounded to fp16 once.  Oracle: the token loop.
#include "yalm_api.h"

#include "gemv.cuh"

#include <cuda_fp16.h>
#include <cuda_runtime.h>

#include <algorithm>
#include <cfloat>
#include <cstdio>

extern "C" int yalm_gemm_f16(float *c, int ldc, const void *a, const void *w, int M, int N, int K, int accumulate, yalm_stream_t s);
extern "C" int yalm_gemm_glu_f16(void *h, int ldh, const void *a, const void *w1, const void *w3, int M, int N, int K, int act,
                                 yalm_stream_t s);

namespace yalm {
namespace prefill {

// x[i, :] = float(E[tokens[i], :])   (src/infer.cu:632-640, for a batch of rows)
__global__ void embed_rows_kernel(float *__restrict__ x, const __half *__restrict__ table, const int *__restrict__ tokens, int dim) {
	const __half *row = table + (size_t)tokens[blockIdx.x] * dim;
	float *dst = x + (size_t)blockIdx.x * dim;
	for (int e = threadIdx.x * 2; e < dim; e += blockDim.x * 2) {
		const float2 v = __half22float2(*reinterpret_cast<const __half2 *>(row + e));
		*reinterpret_cast<float2 *>(dst + e) = v;
	}
}

// out16[i, :] = half(x[i, :] * (1 / sqrt(mean(x^2) + eps)) * w)   (src/infer.cpp:134-144, one CTA per row)
__global__ void __launch_bounds__(256) rmsnorm_rows_kernel(__half *__restrict__ out, const float *__restrict__ x, const float *__restrict__ w,
                                                           int dim, float eps) {
	__shared__ float red[33];
	const float *row = x + (size_t)blockIdx.x * dim;
	float ss = 0.f;
	for (int e = threadIdx.x * 4; e < dim; e += blockDim.x * 4) {
		const float4 v = *reinterpret_cast<const float4 *>(row + e);
		ss += v.x * v.x + v.y * v.y + v.z * v.z + v.w * v.w;
	}
	ss = block_sum(ss, red);
	const float scale = 1.0f / sqrtf(ss / dim + eps);
	__half *dst = out + (size_t)blockIdx.x * dim;
	for (int e = threadIdx.x * 4; e < dim; e += blockDim.x * 4) {
		const float4 v = *reinterpret_cast<const float4 *>(row + e), g = *reinterpret_cast<const float4 *>(w + e);
		*reinterpret_cast<__half2 *>(dst + e) = __floats2half2_rn(v.x * scale * g.x, v.y * scale * g.y);
		*reinterpret_cast<__half2 *>(dst + e + 2) = __floats2half2_rn(v.z * scale * g.z, v.w * scale * g.w);
	}
}

// Row i of qkv (fp32: q | k | v) = position pos0 + i: clip, interleaved-pair RoPE on q and k (src/infer.cpp:200-213,
// 282-292), q -> fp16 (attention operand), k / v -> fp16 rows pos0 + i of the head-major KV ring (src/infer.cpp:294-301).
__global__ void __launch_bounds__(256) rope_append_rows_kernel(__half *__restrict__ q16, __half *__restrict__ kc, __half *__restrict__ vc,
                                                               const float *__restrict__ qkv, int q_dim, int kv_dim, int head_dim,
                                                               int rotary_dim, int max_seq_len, int pos0, float theta, float clip) {
	const int i = blockIdx.x, pos = pos0 + i;
	const float *row = qkv + (size_t)i * (q_dim + 2 * kv_dim);
	for (int p = threadIdx.x; p < (q_dim + 2 * kv_dim) / 2; p += blockDim.x) {
		const int e = 2 * p;
		float v0 = clipf(row[e], clip), v1 = clipf(row[e + 1], clip);
		if (e < q_dim) {
			rope_pair(v0, v1, e % head_dim, pos, theta, rotary_dim);
			*reinterpret_cast<__half2 *>(q16 + (size_t)i * q_dim + e) = __floats2half2_rn(v0, v1);
		} else {
			const bool is_k = e < q_dim + kv_dim;
			const int c = e - q_dim - (is_k ? 0 : kv_dim), h = c / head_dim, j = c % head_dim;
			if (is_k) rope_pair(v0, v1, j, pos, theta, rotary_dim);
			__half *dst = (is_k ? kc : vc) + ((size_t)h * max_seq_len + pos) * head_dim + j;
			*reinterpret_cast<__half2 *>(dst) = __floats2half2_rn(v0, v1);
		}
	}
}

// ---- causal flash attention over the KV ring (no wrap: pos0 + n <= max_seq_len) ----------------------------------------
// grid (ceil(n / 64), n_heads), 4 warps x 16 query rows; K / V tiles of 64 positions in shared memory;
// S = Q K^T and O += P V on mma.sync.m16n8k16 (fp16 in, fp32 accumulate); scores / softmax as src/infer.cpp:216-248:
// s = (q . k) / sqrt(head_dim), p = exp(s - max) / sum.
__device__ __forceinline__ void ldsm4(uint32_t (&r)[4], uint32_t addr) {
	asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0, %1, %2, %3}, [%4];" : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(addr));
}
__device__ __forceinline__ void ldsm4t(uint32_t (&r)[4], uint32_t addr) {
	asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0, %1, %2, %3}, [%4];"
	             : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3])
	             : "r"(addr));
}
__device__ __forceinline__ void mma16816(float (&d)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
	asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0, %1, %2, %3}, {%4, %5, %6, %7}, {%8, %9}, {%0, %1, %2, %3};"
	             : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
	             : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
__device__ __forceinline__ uint32_t pack_h2(float a, float b) {
	const __half2 h = __floats2half2_rn(a, b);
	return *reinterpret_cast<const uint32_t *>(&h);
}

template <int HD>
__global__ void __launch_bounds__(128) attn_prefill_kernel(__half *__restrict__ out, const __half *__restrict__ q16, const __half *__restrict__ kc,
                                                           const __half *__restrict__ vc, int n, int pos0, int n_heads, int n_kv_heads,
                                                           int max_seq_len) {
	constexpr int LD = HD + 8;  // padded row (halves): 16 bytes of skew keep ldmatrix conflict-free
	constexpr int KS = HD / 16; // k-steps of Q K^T
	constexpr int DT = HD / 8;  // 8-wide output tiles of P V
	extern __shared__ __align__(16) __half smem[];
	__half *Qs = smem, *Ks = Qs + 64 * LD, *Vs = Ks + 64 * LD;
	const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
	const int h = blockIdx.y, kvh = h / (n_heads / n_kv_heads), i0 = blockIdx.x * 64;
	const int q_dim = n_heads * HD;
	const __half *kbase = kc + (size_t)kvh * max_seq_len * HD, *vbase = vc + (size_t)kvh * max_seq_len * HD;
	// Q tile (rows past n: zeros)
	for (int c = threadIdx.x; c < 64 * (HD / 8); c += 128) {
		const int r = c / (HD / 8), d = (c % (HD / 8)) * 8;
		uint4 v = make_uint4(0, 0, 0, 0);
		if (i0 + r < n) v = *reinterpret_cast<const uint4 *>(q16 + (size_t)(i0 + r) * q_dim + h * HD + d);
		*reinterpret_cast<uint4 *>(Qs + r * LD + d) = v;
	}
	__syncthreads();
	uint32_t qf[KS][4];
	{
		const int r = warp * 16 + (lane & 7) + ((lane >> 3) & 1) * 8, cofs = (lane >> 4) * 8;
#pragma unroll
		for (int ks = 0; ks < KS; ++ks) ldsm4(qf[ks], (uint32_t)__cvta_generic_to_shared(Qs + r * LD + ks * 16 + cofs));
	}
	float o[DT][4];
#pragma unroll
	for (int t = 0; t < DT; ++t) o[t][0] = o[t][1] = o[t][2] = o[t][3] = 0.f;
	float m0 = -INFINITY, m1 = -INFINITY, l0 = 0.f, l1 = 0.f; // rows lane/4 and lane/4 + 8 of the warp's 16
	const int rA = i0 + warp * 16 + (lane >> 2), rB = rA + 8;
	// causal limits (rows past n reuse the last valid row's limit: finite garbage, never stored)
	const int limA = pos0 + min(rA, n - 1), limB = pos0 + min(rB, n - 1);
	const int last_key = pos0 + min(i0 + 63, n - 1);
	const float div = sqrtf((float)HD);
	for (int j0 = 0; j0 <= last_key; j0 += 64) {
		__syncthreads(); // the previous tile's reads are done
		for (int c = threadIdx.x; c < 64 * (HD / 8); c += 128) {
			const int r = c / (HD / 8), d = (c % (HD / 8)) * 8;
			uint4 kv = make_uint4(0, 0, 0, 0), vv = kv;
			if (j0 + r < max_seq_len && j0 + r <= last_key) {
				kv = *reinterpret_cast<const uint4 *>(kbase + (size_t)(j0 + r) * HD + d);
				vv = *reinterpret_cast<const uint4 *>(vbase + (size_t)(j0 + r) * HD + d);
			}
			*reinterpret_cast<uint4 *>(Ks + r * LD + d) = kv;
			*reinterpret_cast<uint4 *>(Vs + r * LD + d) = vv;
		}
		__syncthreads();
		// ---- S = Q K^T: 8 key tiles of 8 ----
		float s[8][4];
#pragma unroll
		for (int t = 0; t < 8; ++t) s[t][0] = s[t][1] = s[t][2] = s[t][3] = 0.f;
		{
			const int kr = (lane & 7) + (lane >> 4) * 8, kcofs = ((lane >> 3) & 1) * 8;
#pragma unroll
			for (int ks = 0; ks < KS; ++ks)
#pragma unroll
				for (int t = 0; t < 8; t += 2) {
					uint32_t b[4];
					ldsm4(b, (uint32_t)__cvta_generic_to_shared(Ks + (t * 8 + kr) * LD + ks * 16 + kcofs));
					mma16816(s[t], qf[ks], b[0], b[1]);
					mma16816(s[t + 1], qf[ks], b[2], b[3]);
				}
		}
		// ---- scale, causal mask, online softmax (fp32) ----
		float mxA = -INFINITY, mxB = -INFINITY;
#pragma unroll
		for (int t = 0; t < 8; ++t) {
			const int key = j0 + t * 8 + (lane & 3) * 2;
			s[t][0] = key <= limA ? s[t][0] / div : -INFINITY;
			s[t][1] = key + 1 <= limA ? s[t][1] / div : -INFINITY;
			s[t][2] = key <= limB ? s[t][2] / div : -INFINITY;
			s[t][3] = key + 1 <= limB ? s[t][3] / div : -INFINITY;
			mxA = fmaxf(mxA, fmaxf(s[t][0], s[t][1]));
			mxB = fmaxf(mxB, fmaxf(s[t][2], s[t][3]));
		}
		mxA = fmaxf(mxA, __shfl_xor_sync(kFull, mxA, 1)), mxA = fmaxf(mxA, __shfl_xor_sync(kFull, mxA, 2));
		mxB = fmaxf(mxB, __shfl_xor_sync(kFull, mxB, 1)), mxB = fmaxf(mxB, __shfl_xor_sync(kFull, mxB, 2));
		const float nA = fmaxf(m0, mxA), nB = fmaxf(m1, mxB); // finite from the first tile on (key 0 is visible to every row)
		const float cA = m0 == -INFINITY ? 0.f : __expf(m0 - nA), cB = m1 == -INFINITY ? 0.f : __expf(m1 - nB);
		m0 = nA, m1 = nB;
		l0 *= cA, l1 *= cB;
#pragma unroll
		for (int t = 0; t < DT; ++t) o[t][0] *= cA, o[t][1] *= cA, o[t][2] *= cB, o[t][3] *= cB;
		uint32_t pf[4][4]; // P as A fragments: key steps of 16
#pragma unroll
		for (int t = 0; t < 8; ++t) {
			const float p0 = __expf(s[t][0] - nA), p1 = __expf(s[t][1] - nA), p2 = __expf(s[t][2] - nB), p3 = __expf(s[t][3] - nB);
			l0 += p0 + p1, l1 += p2 + p3;
			pf[t >> 1][(t & 1) * 2] = pack_h2(p0, p1);
			pf[t >> 1][(t & 1) * 2 + 1] = pack_h2(p2, p3);
		}
		// ---- O += P V ----
		{
			const int vr = (lane & 7) + ((lane >> 3) & 1) * 8, vcofs = (lane >> 4) * 8;
#pragma unroll
			for (int kk = 0; kk < 4; ++kk)
#pragma unroll
				for (int t = 0; t < DT; t += 2) {
					uint32_t b[4];
					ldsm4t(b, (uint32_t)__cvta_generic_to_shared(Vs + (kk * 16 + vr) * LD + t * 8 + vcofs));
					mma16816(o[t], pf[kk], b[0], b[1]);
					mma16816(o[t + 1], pf[kk], b[2], b[3]);
				}
		}
	}
	l0 += __shfl_xor_sync(kFull, l0, 1), l0 += __shfl_xor_sync(kFull, l0, 2);
	l1 += __shfl_xor_sync(kFull, l1, 1), l1 += __shfl_xor_sync(kFull, l1, 2);
#pragma unroll
	for (int t = 0; t < DT; ++t) {
		const int d = h * HD + t * 8 + (lane & 3) * 2;
		if (rA < n) *reinterpret_cast<__half2 *>(out + (size_t)rA * q_dim + d) = __floats2half2_rn(o[t][0] / l0, o[t][1] / l0);
		if (rB < n) *reinterpret_cast<__half2 *>(out + (size_t)rB * q_dim + d) = __floats2half2_rn(o[t][2] / l1, o[t][3] / l1);
	}
}

template <int HD>
static int launch_attn(__half *out, const __half *q16, const __half *kc, const __half *vc, int n, int pos0, int n_heads, int n_kv_heads,
                       int max_seq_len, cudaStream_t st) {
	const size_t smem = (size_t)3 * 64 * (HD + 8) * sizeof(__half);
	static bool configured = false;
	if (!configured && smem > 48 * 1024) {
		YALM_TRY(cudaFuncSetAttribute(attn_prefill_kernel<HD>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
		configured = true;
	}
	attn_prefill_kernel<HD><<<dim3((n + 63) / 64, n_heads), 128, smem, st>>>(out, q16, kc, vc, n, pos0, n_heads, n_kv_heads, max_seq_len);
	return (int)cudaGetLastError();
}

} // namespace prefill
} // namespace yalm

using namespace yalm;
using namespace yalm::prefill;

// Causal attention of n query rows (positions pos0 .. pos0 + n - 1) over the head-major fp16 KV ring.  Device pointers.
extern "C" int yalm_attn_prefill(void *out16, const void *q16, const uint16_t *key_cache, const uint16_t *value_cache, int n, int pos0,
                                 int n_heads, int n_kv_heads, int head_dim, int max_seq_len, yalm_stream_t s) {
	if (!out16 || !q16 || !key_cache || !value_cache || n < 1 || pos0 < 0 || pos0 + n > max_seq_len || n_kv_heads < 1 || n_heads % n_kv_heads)
		return YALM_EINVAL;
	auto o = static_cast<__half *>(out16);
	auto q = static_cast<const __half *>(q16);
	auto k = reinterpret_cast<const __half *>(key_cache), v = reinterpret_cast<const __half *>(value_cache);
	cudaStream_t st = (cudaStream_t)s;
	switch (head_dim) {
	case 128: return launch_attn<128>(o, q, k, v, n, pos0, n_heads, n_kv_heads, max_seq_len, st);
	case 64: return launch_attn<64>(o, q, k, v, n, pos0, n_heads, n_kv_heads, max_seq_len, st);
	case 32: return launch_attn<32>(o, q, k, v, n, pos0, n_heads, n_kv_heads, max_seq_len, st);
	}
	return YALM_EINVAL;
}

// ---- Model::forward_batch ------------------------------------------------------------------------------------------
struct Model::PrefillState {
	int cap = 0;
	char *arena = nullptr;
	float *x = nullptr, *qkv = nullptr, *logits = nullptr, *probs = nullptr;
	__half *xb = nullptr, *q16 = nullptr, *att = nullptr, *hb = nullptr;
	int *tokens = nullptr, *targets = nullptr;
	~PrefillState() { cudaFree(arena); }
};

bool Model::supports_batch() const {
	const Config &c = *config;
	const char *env = getenv("YALM_B200_PREFILL");
	if (env && env[0] == '0') return false;
	return _device == Device::CUDA && c.weight_dtype == DType::F16 && c.n_experts == 0 && c.tp_size == 1 &&
	       (c.head_dim == 128 || c.head_dim == 64 || c.head_dim == 32) && c.dim % 8 == 0 && c.hidden_dim % 8 == 0 &&
	       (c.n_heads * c.head_dim) % 8 == 0 && (c.n_kv_heads * c.head_dim) % 8 == 0;
}

int Model::batch_capacity() const {
	return supports_batch() ? 512 : 0;
}

int Model::forward_batch(InferenceState &s, const int *tokens, int n, int pos0, const int *targets, float *probs_out) {
	if (!supports_batch() || s.device() != Device::CUDA) return YALM_EINVAL;
	const Config &c = *config;
	const int cap = batch_capacity();
	if (!tokens || n < 1 || n > cap || pos0 < 0 || pos0 + n > c.max_seq_len || (targets && !probs_out)) return YALM_EINVAL;
	for (int i = 0; i < n; ++i)
		if (tokens[i] < 0 || tokens[i] >= c.vocab_size) return YALM_EINVAL;
	const int dim = c.dim, q_dim = c.n_heads * c.head_dim, kv_dim = c.n_kv_heads * c.head_dim, hid = c.hidden_dim, V = c.vocab_size;
	cudaStream_t st = s.stream();
	if (!_prefill) {
		auto ps = std::make_shared<PrefillState>();
		ps->cap = cap;
		auto pad = [](size_t b) { return (b + 255) / 256 * 256; };
		const size_t sz[] = {pad(4ull * cap * dim), pad(4ull * cap * (q_dim + 2 * kv_dim)), pad(4ull * cap * V), pad(4ull * cap),
		                     pad(2ull * cap * std::max(dim, q_dim)), pad(2ull * cap * q_dim), pad(2ull * cap * q_dim), pad(2ull * cap * hid),
		                     pad(4ull * cap), pad(4ull * cap)};
		size_t total = 0;
		for (size_t b : sz) total += b;
		if (cudaMalloc((void **)&ps->arena, total) != cudaSuccess) {
			cudaGetLastError();
			return (int)cudaErrorMemoryAllocation;
		}
		char *p = ps->arena;
		void **slots[] = {(void **)&ps->x, (void **)&ps->qkv, (void **)&ps->logits, (void **)&ps->probs, (void **)&ps->xb,
		                  (void **)&ps->q16, (void **)&ps->att, (void **)&ps->hb, (void **)&ps->tokens, (void **)&ps->targets};
		for (int i = 0; i < 10; ++i) *slots[i] = p, p += sz[i];
		_prefill = ps;
	}
	PrefillState &P = *_prefill;
	YALM_TRY(cudaMemcpyAsync(P.tokens, tokens, sizeof(int) * n, cudaMemcpyHostToDevice, st));
	if (targets) YALM_TRY(cudaMemcpyAsync(P.targets, targets, sizeof(int) * n, cudaMemcpyHostToDevice, st));
	embed_rows_kernel<<<n, 256, 0, st>>>(P.x, static_cast<const __half *>(token_embedding_table), P.tokens, dim);
	const int act = c.act == ActivationType::SILU ? YALM_ACT_SILU : YALM_ACT_GELU;
	int rc;
	for (auto &bp : blocks) {
		const Block &b = *bp;
		rmsnorm_rows_kernel<<<n, 256, 0, st>>>(P.xb, P.x, b._rms_att_weight, dim, c.norm_eps);
		// q | k | v of the chunk: three GEMMs into one (n, q_dim + 2 kv_dim) fp32 buffer
		const int ld = q_dim + 2 * kv_dim;
		if ((rc = yalm_gemm_f16(P.qkv, ld, P.xb, b._wq, n, q_dim, dim, 0, st))) return rc;
		if ((rc = yalm_gemm_f16(P.qkv + q_dim, ld, P.xb, b._wk, n, kv_dim, dim, 0, st))) return rc;
		if ((rc = yalm_gemm_f16(P.qkv + q_dim + kv_dim, ld, P.xb, b._wv, n, kv_dim, dim, 0, st))) return rc;
		rope_append_rows_kernel<<<n, 256, 0, st>>>(P.q16, reinterpret_cast<__half *>(b._key_cache), reinterpret_cast<__half *>(b._value_cache),
		                                          P.qkv, q_dim, kv_dim, c.head_dim, c.rotary_dim, c.max_seq_len, pos0, c.rope_theta, c.qkv_clip);
		if ((rc = yalm_attn_prefill(P.att, P.q16, b._key_cache, b._value_cache, n, pos0, c.n_heads, c.n_kv_heads, c.head_dim, c.max_seq_len, st)))
			return rc;
		if ((rc = yalm_gemm_f16(P.x, dim, P.att, b._wo, n, dim, q_dim, 1, st))) return rc; // x += Wo att
		rmsnorm_rows_kernel<<<n, 256, 0, st>>>(P.xb, P.x, b._rms_ffn_weight, dim, c.norm_eps);
		if ((rc = yalm_gemm_glu_f16(P.hb, hid, P.xb, b._w1, b._w3, n, hid, dim, act, st))) return rc;
		if ((rc = yalm_gemm_f16(P.x, dim, P.hb, b._w2, n, dim, hid, 1, st))) return rc; // x += W2 hb
	}
	if (targets) { // per-position logits -> probability of the given next token (Sampler::sample_prob per row)
		rmsnorm_rows_kernel<<<n, 256, 0, st>>>(P.xb, P.x, rms_final_weight, dim, c.norm_eps);
		if ((rc = yalm_gemm_f16(P.logits, V, P.xb, wcls, n, V, dim, 0, st))) return rc;
		if ((rc = yalm_token_probs(P.logits, V, n, P.targets, P.probs, st))) return rc;
		YALM_TRY(cudaMemcpyAsync(probs_out, P.probs, sizeof(float) * n, cudaMemcpyDeviceToHost, st));
	}
	// the LAST position's logits through the decode path's final norm + LM head (fp32 activations), as after forward()
	if ((rc = yalm_lm_head(s.logits_dev(), P.x + (size_t)(n - 1) * dim, rms_final_weight, wcls, wcls_scale, dim, V, c.norm_eps,
	                       static_cast<int>(c.weight_dtype), st)))
		return rc;
	if (s.host_logits())
		YALM_TRY(cudaMemcpyAsync(s.logits(), s.logits_dev(), sizeof(float) * V, cudaMemcpyDeviceToHost, st));
	YALM_TRY(cudaStreamSynchronize(st));
	return (int)cudaGetLastError();
}
