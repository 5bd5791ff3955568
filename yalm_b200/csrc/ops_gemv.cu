// GEMV-family decode ops for sm_100a: embed, rmsnorm+QKV+clip+RoPE+KV-append, GEMV+residual (dense
// and MoE mix), rmsnorm+[router]+w1/w3+GLU, final rmsnorm+LM head, argmax, bare GEMV.
// C ABI in include/yalm_b200.h; each launcher cites the reference kernels it replaces there.
#include "gemv.cuh"
#include "launch.h"

#include <cfloat>

namespace yalm {

constexpr int kThreads = 256; // 8 warps per CTA

#define YALM_DISPATCH_DT(dt, ...)                                                                             \
	switch (dt) {                                                                                             \
	case YALM_F32: { constexpr int DT = YALM_F32; __VA_ARGS__; break; }                                       \
	case YALM_F16: { constexpr int DT = YALM_F16; __VA_ARGS__; break; }                                       \
	case YALM_F8E4M3: { constexpr int DT = YALM_F8E4M3; __VA_ARGS__; break; }                                 \
	default: return YALM_EINVAL;                                                                              \
	}

// ------------------------------------------------------------------------------------------------
// embed + ring position
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ void derive_ring(int *ctl, int pos, int max_seq_len) {
	// src/infer.cu:1081-1083
	int kv_sink = pos >= max_seq_len ? KV_SINKS : 0;
	ctl[CTL_SINK] = kv_sink;
	ctl[CTL_KVPOS] = kv_sink + (pos - kv_sink) % (max_seq_len - kv_sink);
	ctl[CTL_KVLEN] = pos >= max_seq_len ? max_seq_len : pos + 1;
}

template <int DT>
__global__ void embed_kernel(float *__restrict__ x, const void *__restrict__ table, const float *__restrict__ row_scale,
                             int dim, int *ctl, int token_arg, int pos_arg, int max_seq_len) {
	int token = token_arg, pos = pos_arg;
	if (token_arg < 0) {
		token = ctl[CTL_TOKEN];
		pos = ctl[CTL_POS];
	}
	const float sc = (DT == YALM_F8E4M3 && row_scale) ? row_scale[token] : 1.0f;
	for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < dim; i += gridDim.x * blockDim.x)
		x[i] = WTraits<DT>::to_float(table, (size_t)token * dim + i) * sc;
	if (blockIdx.x == 0 && threadIdx.x == 0 && token_arg >= 0) {
		ctl[CTL_TOKEN] = token;
		ctl[CTL_POS] = pos;
		derive_ring(ctl, pos, max_seq_len);
	}
}

// ------------------------------------------------------------------------------------------------
// rmsnorm -> QKV -> clip -> RoPE -> KV append (+ sink re-rotation)
// ------------------------------------------------------------------------------------------------
struct QkvArgs {
	float *q_out;
	__half *kb, *vb;
	const float *x, *rms_w;
	const char *wq, *wk, *wv;
	const float *sq, *sk, *sv;
	int dim, q_dim, kv_dim, head_dim, rotary_dim, max_seq_len;
	float theta, eps, clip;
	const int *ctl;
};

template <int DT, int U>
__global__ void __launch_bounds__(kThreads) qkv_rope_append_kernel(QkvArgs a) {
	extern __shared__ __align__(16) float smem[];
	float *red = smem;      // 36 floats
	float *xs = smem + 36;
	constexpr int EP = WTraits<DT>::EP, B = WTraits<DT>::BYTES;
	stage_x_rmsnorm<EP>(xs, a.x, a.rms_w, a.dim, a.eps, red);
	__syncthreads();

	const int pos = a.ctl[CTL_POS], kv_pos = a.ctl[CTL_KVPOS], kv_sink = a.ctl[CTL_SINK];
	const int lane = threadIdx.x & 31;
	const int warps = (gridDim.x * blockDim.x) >> 5, gw = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
	const int ntasks = (a.q_dim + 2 * a.kv_dim) >> 1;
	const size_t row_bytes = (size_t)a.dim * B;
	for (int t = gw; t < ntasks; t += warps) {
		int row = 2 * t, which = 0; // 0 q, 1 k, 2 v
		const char *w = a.wq;
		const float *sc = a.sq;
		if (row >= a.q_dim + a.kv_dim) {
			which = 2, row -= a.q_dim + a.kv_dim, w = a.wv, sc = a.sv;
		} else if (row >= a.q_dim) {
			which = 1, row -= a.q_dim, w = a.wk, sc = a.sk;
		}
		const char *rows[2] = {w + (size_t)row * row_bytes, w + (size_t)(row + 1) * row_bytes};
		float acc[2];
		warp_rows_dot<DT, 2, U>(rows, xs, a.dim, acc);
		if (lane == 0) {
			float v0 = acc[0], v1 = acc[1];
			if (DT == YALM_F8E4M3 && sc) v0 *= sc[row], v1 *= sc[row + 1];
			v0 = clipf(v0, a.clip), v1 = clipf(v1, a.clip);
			if (which < 2) rope_pair(v0, v1, row % a.head_dim, pos, a.theta, a.rotary_dim);
			if (which == 0) {
				*reinterpret_cast<float2 *>(a.q_out + row) = make_float2(v0, v1);
			} else {
				// head-major KV ring: (n_kv_heads, max_seq_len, head_dim)
				__half *dst = (which == 1 ? a.kb : a.vb) + ((size_t)(row / a.head_dim) * a.max_seq_len + kv_pos) * a.head_dim +
				              row % a.head_dim;
				*reinterpret_cast<__half2 *>(dst) = __floats2half2_rn(v0, v1);
			}
		}
	}
	// Attention sinks: rotate each sink key forward by one position, fp16 -> fp32 -> fp16
	// (src/infer.cpp:303-317, src/infer.cu:679-697).  Rows 0..kv_sink-1 never collide with kv_pos.
	if (kv_sink > 0 && blockIdx.x == gridDim.x - 1) {
		const int npairs = kv_sink * (a.kv_dim >> 1);
		for (int p = threadIdx.x; p < npairs; p += blockDim.x) {
			const int r = p / (a.kv_dim >> 1), i = (p % (a.kv_dim >> 1)) * 2;
			__half2 *ptr = reinterpret_cast<__half2 *>(a.kb + ((size_t)(i / a.head_dim) * a.max_seq_len + r) * a.head_dim +
			                                           i % a.head_dim);
			float2 v = __half22float2(*ptr);
			rope_pair(v.x, v.y, i % a.head_dim, 1, a.theta, a.rotary_dim);
			*ptr = __floats2half2_rn(v.x, v.y);
		}
	}
}

// ------------------------------------------------------------------------------------------------
// out = residual + sum_k aw[k] * W[e_k] v_k
// ------------------------------------------------------------------------------------------------
struct ResidArgs {
	float *out;
	const float *residual, *v;
	const char *w;
	const float *scale;
	int n, d, n_active;
	const int *active_e;
	const float *active_w;
	int ep_rank, ep_size; // expert parallelism: this rank holds experts e with e % ep_size == ep_rank, stacked in order (ep_size 1: all)
};

template <int DT, int R, int U>
__global__ void __launch_bounds__(kThreads) gemv_residual_kernel(ResidArgs a) {
	extern __shared__ __align__(16) float smem[];
	constexpr int EP = WTraits<DT>::EP, B = WTraits<DT>::BYTES;
	const int npad = xs_padded<EP>(a.n);
	for (int k = 0; k < a.n_active; ++k) stage_x_plain<EP>(smem + k * npad, a.v + (size_t)k * a.n, a.n);
	__syncthreads();
	const int lane = threadIdx.x & 31;
	const int warps = (gridDim.x * blockDim.x) >> 5, gw = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
	const int ntasks = a.d / R;
	const size_t row_bytes = (size_t)a.n * B;
	for (int t = gw; t < ntasks; t += warps) {
		const int row = t * R;
		float res[R];
#pragma unroll
		for (int r = 0; r < R; ++r) res[r] = a.residual ? a.residual[row + r] : 0.f;
		for (int k = 0; k < a.n_active; ++k) {
			int e = a.active_e ? a.active_e[k] : 0;
			const float ew = a.active_w ? a.active_w[k] : 1.0f;
			if (a.ep_size > 1) { // another rank's expert: its contribution arrives through the all-reduce
				if (e % a.ep_size != a.ep_rank) continue;
				e /= a.ep_size;
			}
			const char *base = a.w + ((size_t)e * a.d + row) * row_bytes;
			const char *rows[R];
#pragma unroll
			for (int r = 0; r < R; ++r) rows[r] = base + r * row_bytes;
			float acc[R];
			warp_rows_dot<DT, R, U>(rows, smem + k * npad, a.n, acc);
#pragma unroll
			for (int r = 0; r < R; ++r) {
				float v = acc[r];
				if (DT == YALM_F8E4M3 && a.scale) v *= a.scale[(size_t)e * a.d + row + r];
				// reference: x += xb2 * expert_weight (src/infer.cpp:381-383); dense: x += hb (:335-337)
				res[r] = a.active_w ? res[r] + v * ew : res[r] + v;
			}
		}
		if (lane < R) {
#pragma unroll
			for (int r = 0; r < R; ++r)
				if (lane == r) a.out[row + r] = res[r];
		}
	}
}

// ------------------------------------------------------------------------------------------------
// rmsnorm -> [router + top-k] -> hb = act(W1 xb) * (W3 xb)
// ------------------------------------------------------------------------------------------------
struct FfnArgs {
	float *hb;
	const float *x, *rms_w;
	const char *w1, *w3, *gate;
	const float *s1, *s3, *sgate;
	int dim, hidden, act;
	float eps;
	int n_experts, n_active;
	int *active_e;
	float *active_w, *moe_logits;
	int ep_rank, ep_size; // expert parallelism (see ResidArgs)
};

constexpr int kMaxExperts = 64; // moe_gate's uint64_t mask (src/infer.cpp:111)

template <int DT, int P, int U>
__global__ void __launch_bounds__(kThreads) ffn_glu_kernel(FfnArgs a) {
	extern __shared__ __align__(16) float smem[];
	float *red = smem;                   // 36
	float *gl = smem + 36;               // kMaxExperts router logits
	float *sel_w = gl + kMaxExperts;     // kMaxExperts
	int *sel_e = reinterpret_cast<int *>(sel_w + kMaxExperts); // kMaxExperts
	float *xs = smem + 36 + 3 * kMaxExperts;
	constexpr int EP = WTraits<DT>::EP, B = WTraits<DT>::BYTES;
	if (a.rms_w) stage_x_rmsnorm<EP>(xs, a.x, a.rms_w, a.dim, a.eps, red);
	else stage_x_plain<EP>(xs, a.x, a.dim); // test entry ffn_cuda: no norm in front
	__syncthreads();
	const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5, nw = blockDim.x >> 5;
	const size_t row_bytes = (size_t)a.dim * B;

	int n_active = 1;
	if (a.n_experts > 0) {
		// Router: every CTA recomputes the n_experts logits (n_experts*dim weights from L2) and the
		// top-k, bit-identically, so expert ids never leave the device and no extra launch is needed.
		for (int e = wid; e < a.n_experts; e += nw) {
			const char *rows[1] = {a.gate + (size_t)e * row_bytes};
			float acc[1];
			warp_rows_dot<DT, 1, 4>(rows, xs, a.dim, acc);
			if (lane == 0) gl[e] = (DT == YALM_F8E4M3 && a.sgate) ? acc[0] * a.sgate[e] : acc[0];
		}
		__syncthreads();
		if (threadIdx.x == 0) {
			// moe_gate, src/infer.cpp:100-132: strict '>' arg-max (lowest index wins ties), weights =
			// softmax over the selected experts only, exponentials relative to the global max.
			float max_val = -FLT_MAX;
			for (int j = 0; j < a.n_experts; ++j)
				if (gl[j] > max_val) max_val = gl[j];
			unsigned long long mask = 0;
			float wsum = 0.f;
			for (int k = 0; k < a.n_active; ++k) {
				int best = -1;
				for (int j = 0; j < a.n_experts; ++j)
					if ((mask & (1ull << j)) == 0 && (best == -1 || gl[j] > gl[best])) best = j;
				sel_e[k] = best;
				wsum += expf(gl[best] - max_val);
				mask |= 1ull << best;
			}
			for (int k = 0; k < a.n_active; ++k) sel_w[k] = expf(gl[sel_e[k]] - max_val) / wsum;
		}
		__syncthreads();
		n_active = a.n_active;
		if (blockIdx.x == 0 && threadIdx.x < n_active) {
			a.active_e[threadIdx.x] = sel_e[threadIdx.x];
			a.active_w[threadIdx.x] = sel_w[threadIdx.x];
		}
		if (blockIdx.x == 0 && a.moe_logits && threadIdx.x < a.n_experts) a.moe_logits[threadIdx.x] = gl[threadIdx.x];
	}

	const int warps = (gridDim.x * blockDim.x) >> 5, gw = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
	const int per_expert = a.hidden / P;
	const int ntasks = n_active * per_expert;
	for (int t = gw; t < ntasks; t += warps) {
		const int k = t / per_expert, i0 = (t % per_expert) * P;
		int e = a.n_experts > 0 ? sel_e[k] : 0;
		if (a.ep_size > 1) { // only the rank that holds expert e computes its hb rows
			if (e % a.ep_size != a.ep_rank) continue;
			e /= a.ep_size;
		}
		const size_t r0 = (size_t)e * a.hidden + i0;
		const char *rows[2 * P];
#pragma unroll
		for (int p = 0; p < P; ++p) {
			rows[2 * p] = a.w1 + (r0 + p) * row_bytes;
			rows[2 * p + 1] = a.w3 + (r0 + p) * row_bytes;
		}
		float acc[2 * P];
		warp_rows_dot<DT, 2 * P, U>(rows, xs, a.dim, acc);
#pragma unroll
		for (int p = 0; p < P; ++p) {
			if (lane == p) {
				float g = acc[2 * p], u = acc[2 * p + 1];
				if (DT == YALM_F8E4M3 && a.s1) g *= a.s1[r0 + p], u *= a.s3[r0 + p];
				g = a.act == YALM_ACT_GELU ? act_gelu(g) : act_silu(g);
				a.hb[(size_t)k * a.hidden + i0 + p] = g * u;
			}
		}
	}
}

// ------------------------------------------------------------------------------------------------
// [rmsnorm ->] plain rows: LM head and the bare GEMV
// ------------------------------------------------------------------------------------------------
struct RowsArgs {
	float *out;
	const float *x, *rms_w; // rms_w == nullptr: no norm
	const char *w;
	const float *scale;
	int n, d;
	float eps;
};

template <int DT, int R, int U>
__global__ void __launch_bounds__(kThreads) gemv_rows_kernel(RowsArgs a) {
	extern __shared__ __align__(16) float smem[];
	float *red = smem, *xs = smem + 36;
	constexpr int EP = WTraits<DT>::EP, B = WTraits<DT>::BYTES;
	if (a.rms_w) stage_x_rmsnorm<EP>(xs, a.x, a.rms_w, a.n, a.eps, red);
	else stage_x_plain<EP>(xs, a.x, a.n);
	__syncthreads();
	const int lane = threadIdx.x & 31;
	const int warps = (gridDim.x * blockDim.x) >> 5, gw = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
	const size_t row_bytes = (size_t)a.n * B;
	const int ntasks = (a.d + R - 1) / R;
	for (int t = gw; t < ntasks; t += warps) {
		const int row = t * R;
		const char *rows[R];
#pragma unroll
		for (int r = 0; r < R; ++r) rows[r] = a.w + (size_t)min(row + r, a.d - 1) * row_bytes;
		float acc[R];
		warp_rows_dot<DT, R, U>(rows, xs, a.n, acc);
#pragma unroll
		for (int r = 0; r < R; ++r)
			if (lane == r && row + r < a.d)
				a.out[row + r] = (DT == YALM_F8E4M3 && a.scale) ? acc[r] * a.scale[row + r] : acc[r];
	}
}

// ------------------------------------------------------------------------------------------------
// greedy arg-max with the sampler's first-max rule, then advance the decode control block
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(1024) argmax_advance_kernel(const float *__restrict__ logits, int vocab, int *ctl,
                                                              int max_seq_len, int *tokens_out) {
	__shared__ float sv[32];
	__shared__ int si[32];
	float best = -FLT_MAX;
	int bi = 0x7fffffff;
	// Sampler::sample_argmax (src/sampler.cpp:27-38): max_val starts at -FLT_MAX, strict '>', so the
	// winner is the LOWEST index among equal maxima, and index 0 if nothing exceeds -FLT_MAX.
	for (int i = threadIdx.x; i < vocab; i += blockDim.x) {
		float v = logits[i];
		if (v > best) best = v, bi = i;
	}
	for (int o = 16; o > 0; o >>= 1) {
		float ov = __shfl_xor_sync(kFull, best, o);
		int oi = __shfl_xor_sync(kFull, bi, o);
		if (ov > best || (ov == best && oi < bi)) best = ov, bi = oi;
	}
	if ((threadIdx.x & 31) == 0) sv[threadIdx.x >> 5] = best, si[threadIdx.x >> 5] = bi;
	__syncthreads();
	if (threadIdx.x < 32) {
		const int nw = blockDim.x >> 5;
		best = threadIdx.x < nw ? sv[threadIdx.x] : -FLT_MAX;
		bi = threadIdx.x < nw ? si[threadIdx.x] : 0x7fffffff;
		for (int o = 16; o > 0; o >>= 1) {
			float ov = __shfl_xor_sync(kFull, best, o);
			int oi = __shfl_xor_sync(kFull, bi, o);
			if (ov > best || (ov == best && oi < bi)) best = ov, bi = oi;
		}
		if (threadIdx.x == 0) {
			if (bi == 0x7fffffff) bi = 0;
			const int step = ctl[CTL_STEP];
			if (tokens_out) tokens_out[step] = bi;
			ctl[CTL_STEP] = step + 1;
			ctl[CTL_TOKEN] = bi;
			const int pos = ctl[CTL_POS] + 1;
			ctl[CTL_POS] = pos;
			derive_ring(ctl, pos, max_seq_len);
		}
	}
}

} // namespace yalm

// ================================================================================================
// C ABI launchers
// ================================================================================================
using namespace yalm;

static inline int ep_of(int dtype) {
	return dtype == YALM_F32 ? 4 : dtype == YALM_F16 ? 8 : dtype == YALM_F8E4M3 ? 16 : 0;
}
static inline size_t xs_bytes(int dtype, int n) {
	const int ch = 32 * ep_of(dtype);
	return (size_t)((n + ch - 1) / ch * ch) * sizeof(float);
}

template <typename K, typename A>
static int launch_persistent(K kernel, const A &args, size_t smem, long tasks, cudaStream_t st) {
	cudaError_t err;
	if ((int)smem > device_info().max_smem_optin) return YALM_EINVAL;
	int grid = persistent_grid((const void *)kernel, kThreads, smem, tasks, &err);
	if (err != cudaSuccess) return (int)err;
	kernel<<<grid, kThreads, smem, st>>>(args);
	return (int)cudaGetLastError();
}

extern "C" int yalm_embed(float *x, const void *table, const float *row_scale, int dim, int dtype, int *ctl,
                          int token, int pos, int max_seq_len, yalm_stream_t s) {
	if (!x || !table || !ctl || dim <= 0 || max_seq_len <= KV_SINKS) return YALM_EINVAL;
	cudaStream_t st = (cudaStream_t)s;
	const int grid = std::max(1, std::min(8, (dim + 1023) / 1024));
	YALM_DISPATCH_DT(dtype, (embed_kernel<DT><<<grid, 1024, 0, st>>>(x, table, row_scale, dim, ctl, token, pos,
	                                                                  max_seq_len)));
	return (int)cudaGetLastError();
}

extern "C" int yalm_qkv_rope_append(float *q_out, uint16_t *key_cache, uint16_t *value_cache, const float *x,
                                    const float *rms_weight, const void *wq, const void *wk, const void *wv,
                                    const float *sq, const float *sk, const float *sv, int dim, int n_heads,
                                    int n_kv_heads, int head_dim, int rotary_dim, int max_seq_len, float rope_theta,
                                    float norm_eps, float qkv_clip, int dtype, const int *ctl, yalm_stream_t s) {
	const int ep = ep_of(dtype);
	if (!ep || !q_out || !key_cache || !value_cache || !x || !rms_weight || !wq || !wk || !wv || !ctl)
		return YALM_EINVAL;
	if (dim % ep || dim % 4 || head_dim % 2 || n_heads <= 0 || n_kv_heads <= 0 || max_seq_len <= 0) return YALM_EINVAL;
	QkvArgs a;
	a.q_out = q_out, a.kb = (__half *)key_cache, a.vb = (__half *)value_cache;
	a.x = x, a.rms_w = rms_weight;
	a.wq = (const char *)wq, a.wk = (const char *)wk, a.wv = (const char *)wv;
	a.sq = sq, a.sk = sk, a.sv = sv;
	a.dim = dim, a.q_dim = n_heads * head_dim, a.kv_dim = n_kv_heads * head_dim;
	a.head_dim = head_dim, a.rotary_dim = rotary_dim, a.max_seq_len = max_seq_len, a.theta = rope_theta, a.eps = norm_eps, a.clip = qkv_clip;
	a.ctl = ctl;
	const size_t smem = 36 * sizeof(float) + xs_bytes(dtype, dim);
	const long tasks = (a.q_dim + 2 * a.kv_dim) / 2;
	YALM_DISPATCH_DT(dtype, return launch_persistent(qkv_rope_append_kernel<DT, 8>, a, smem, tasks, (cudaStream_t)s));
	return 0;
}

extern "C" int yalm_gemv_residual(float *out, const float *residual, const float *v, const void *w,
                                  const float *scale, int n, int d, int dtype, int n_active, const int *active_e,
                                  const float *active_w, yalm_stream_t s) {
	return yalm_gemv_residual_ep(out, residual, v, w, scale, n, d, dtype, n_active, active_e, active_w, 0, 1, s);
}

extern "C" int yalm_gemv_residual_ep(float *out, const float *residual, const float *v, const void *w, const float *scale, int n,
                                     int d, int dtype, int n_active, const int *active_e, const float *active_w, int ep_rank,
                                     int ep_size, yalm_stream_t s) {
	const int ep = ep_of(dtype);
	if (!ep || !out || !v || !w || n <= 0 || d <= 0 || n_active < 1 || ep_size < 1 || ep_rank < 0 || ep_rank >= ep_size) return YALM_EINVAL;
	if (n % ep || d % 2 || (n_active > 1 && (!active_e || !active_w))) return YALM_EINVAL;
	ResidArgs a;
	a.out = out, a.residual = residual, a.v = v, a.w = (const char *)w, a.scale = scale;
	a.n = n, a.d = d, a.n_active = n_active, a.active_e = active_e, a.active_w = active_w;
	a.ep_rank = ep_rank, a.ep_size = ep_size;
	const size_t smem = (size_t)n_active * xs_bytes(dtype, n);
	YALM_DISPATCH_DT(dtype, return launch_persistent(gemv_residual_kernel<DT, 2, 8>, a, smem, d / 2, (cudaStream_t)s));
	return 0;
}

extern "C" int yalm_ffn_glu(float *hb, const float *x, const float *rms_weight, const void *w1, const void *w3,
                            const float *s1, const float *s3, const void *gate, const float *sgate, int dim,
                            int hidden_dim, int act, float norm_eps, int dtype, int n_experts, int n_active,
                            int *active_e, float *active_w, float *moe_logits, yalm_stream_t s) {
	return yalm_ffn_glu_ep(hb, x, rms_weight, w1, w3, s1, s3, gate, sgate, dim, hidden_dim, act, norm_eps, dtype, n_experts, n_active,
	                       active_e, active_w, moe_logits, 0, 1, s);
}

extern "C" int yalm_ffn_glu_ep(float *hb, const float *x, const float *rms_weight, const void *w1, const void *w3, const float *s1,
                               const float *s3, const void *gate, const float *sgate, int dim, int hidden_dim, int act, float norm_eps,
                               int dtype, int n_experts, int n_active, int *active_e, float *active_w, float *moe_logits, int ep_rank,
                               int ep_size, yalm_stream_t s) {
	if (ep_size < 1 || ep_rank < 0 || ep_rank >= ep_size || (ep_size > 1 && (n_experts < 1 || n_experts % ep_size))) return YALM_EINVAL;
	const int ep = ep_of(dtype);
	if (!ep || !hb || !x || !w1 || !w3 || dim % ep || dim % 4 || hidden_dim <= 0) return YALM_EINVAL;
	if (n_experts < 0 || n_experts > kMaxExperts) return YALM_EINVAL;
	if (n_experts > 0 && (!gate || !active_e || !active_w || n_active < 1 || n_active > n_experts)) return YALM_EINVAL;
	FfnArgs a;
	a.hb = hb, a.x = x, a.rms_w = rms_weight;
	a.w1 = (const char *)w1, a.w3 = (const char *)w3, a.gate = (const char *)gate;
	a.s1 = s1, a.s3 = s3, a.sgate = sgate;
	a.dim = dim, a.hidden = hidden_dim, a.act = act, a.eps = norm_eps;
	a.n_experts = n_experts, a.n_active = n_experts > 0 ? n_active : 1;
	a.active_e = active_e, a.active_w = active_w, a.moe_logits = moe_logits;
	a.ep_rank = ep_rank, a.ep_size = ep_size;
	const size_t smem = (36 + 3 * kMaxExperts) * sizeof(float) + xs_bytes(dtype, dim);
	const long k = a.n_active;
	if (hidden_dim % 2 == 0) {
		YALM_DISPATCH_DT(dtype, return launch_persistent(ffn_glu_kernel<DT, 2, 4>, a, smem, k * hidden_dim / 2,
		                                                  (cudaStream_t)s));
	} else {
		YALM_DISPATCH_DT(dtype, return launch_persistent(ffn_glu_kernel<DT, 1, 8>, a, smem, k * hidden_dim,
		                                                  (cudaStream_t)s));
	}
	return 0;
}

static int rows_launch(float *out, const float *x, const float *rms_w, const void *w, const float *scale, int n,
                       int d, float eps, int dtype, cudaStream_t st) {
	const int ep = ep_of(dtype);
	if (!ep || !out || !x || !w || n <= 0 || d <= 0 || n % ep) return YALM_EINVAL;
	RowsArgs a;
	a.out = out, a.x = x, a.rms_w = rms_w, a.w = (const char *)w, a.scale = scale, a.n = n, a.d = d, a.eps = eps;
	const size_t smem = 36 * sizeof(float) + xs_bytes(dtype, n);
	YALM_DISPATCH_DT(dtype, return launch_persistent(gemv_rows_kernel<DT, 4, 4>, a, smem, (d + 3) / 4, st));
	return 0;
}

extern "C" int yalm_lm_head(float *logits, const float *x, const float *rms_weight, const void *wcls,
                            const float *scale, int dim, int vocab, float norm_eps, int dtype, yalm_stream_t s) {
	if (!rms_weight) return YALM_EINVAL;
	return rows_launch(logits, x, rms_weight, wcls, scale, dim, vocab, norm_eps, dtype, (cudaStream_t)s);
}

extern "C" int yalm_gemv(float *xout, const float *x, const void *w, const float *scale, int n, int d, int dtype,
                         yalm_stream_t s) {
	return rows_launch(xout, x, nullptr, w, scale, n, d, 0.f, dtype, (cudaStream_t)s);
}

extern "C" int yalm_argmax_advance(const float *logits, int vocab, int *ctl, int max_seq_len, int *tokens_out,
                                   yalm_stream_t s) {
	if (!logits || !ctl || vocab <= 0 || max_seq_len <= KV_SINKS) return YALM_EINVAL;
	argmax_advance_kernel<<<1, 1024, 0, (cudaStream_t)s>>>(logits, vocab, ctl, max_seq_len, tokens_out);
	return (int)cudaGetLastError();
}
