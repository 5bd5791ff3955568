// Split-K flash-decode GQA attention over the fp16 KV ring buffer (sm_100a).
//
// Replaces the reference's three kernels attn_dot / attn_softmax / att_mix (src/infer.cu:338-524):
//  * one CTA serves ALL q-heads of a KV head, so every K/V row is read once, not n_heads/n_kv_heads
//    times, and each row is read as 16-byte lane-contiguous pieces (the reference walks one 256-byte
//    key row per thread);
//  * the sequence is split over gridDim.x CTAs per KV head (the reference's att_mix runs 32 CTAs on a
//    148-SM part) with an online softmax per split; the last CTA of each KV head to finish merges the
//    split partials (arrival counter in the workspace) — no second launch;
//  * no (n_heads, max_seq_len) score matrix round-trips through memory.
// Softmax order is irrelevant to the ring buffer, so sliding-window positions and the two attention
// sinks (rows 0-1) need no unrolling: the kernel just attends over rows [0, kv_len).
//
// Arithmetic follows the CPU reference (src/infer.cpp:216-248): fp32 dot over fp16 K, score /
// sqrtf(head_dim), exp(x - max) / sum, fp32 mix over fp16 V.
#include "common.cuh"
#include "launch.h"

#include <cfloat>
#include <cstdio>
#include <mutex>

namespace yalm {

constexpr int kAttnThreads = 256;
constexpr int kAttnWarps = kAttnThreads / 32;
constexpr int kTile = 512; // rows per online-softmax tile

struct AttnArgs {
	float *out;            // (n_heads*head_dim)
	const __half *kb, *vb; // head-major ring: (n_kv_heads, max_seq_len, head_dim)
	const float *q;        // (n_heads*head_dim)
	float *ws_acc;         // (n_kv_heads, S, G, HD) unnormalised partial outputs
	float *ws_ml;          // (n_kv_heads, S, G, 2)  running max, running sum
	unsigned *counters;    // (n_kv_heads) arrival counters, zero between launches
	int n_kv_heads, max_seq_len;
	const int *ctl;
};

__device__ __forceinline__ void unpack8(uint4 v, float (&f)[8]) {
	float2 a = __half22float2(*reinterpret_cast<const __half2 *>(&v.x));
	float2 b = __half22float2(*reinterpret_cast<const __half2 *>(&v.y));
	float2 c = __half22float2(*reinterpret_cast<const __half2 *>(&v.z));
	float2 d = __half22float2(*reinterpret_cast<const __half2 *>(&v.w));
	f[0] = a.x, f[1] = a.y, f[2] = b.x, f[3] = b.y, f[4] = c.x, f[5] = c.y, f[6] = d.x, f[7] = d.y;
}

template <int HD, int G>
__global__ void __launch_bounds__(kAttnThreads) attn_decode_kernel(AttnArgs a) {
	constexpr int L = HD / 8;    // lanes per K/V row (16 bytes each)
	constexpr int RPW = 32 / L;  // rows per warp step
	constexpr int RPC = RPW * kAttnWarps; // rows per CTA step
	constexpr int UN = 4;        // row-steps in flight per thread

	extern __shared__ __align__(16) float smem[];
	float *red = smem;                       // 36
	float *p_s = smem + 36;                  // [kTile][G]
	float *o_s = p_s + kTile * G;            // [kAttnWarps][G][HD]
	__shared__ float m_run[G], l_run[G], corr_s[G];
	__shared__ unsigned is_last;

	const int kvh = blockIdx.y, S = gridDim.x, split = blockIdx.x;
	const int kv_len = a.ctl[CTL_KVLEN];
	constexpr int kv_dim = HD; // row stride inside one head's slab
	const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
	const int sub = lane % L, rgrp = lane / L;

	// this CTA's row range
	int per = (kv_len + S - 1) / S;
	per = (per + RPC - 1) / RPC * RPC;
	const int t_begin = min(split * per, kv_len), t_end = min(t_begin + per, kv_len);

	// q slice of this lane for each of the G heads: dims [sub*8, sub*8+8)
	float qr[G][8];
#pragma unroll
	for (int g = 0; g < G; ++g) {
		const float *qp = a.q + (size_t)(kvh * G + g) * HD + sub * 8;
		float4 lo = *reinterpret_cast<const float4 *>(qp), hi = *reinterpret_cast<const float4 *>(qp + 4);
		qr[g][0] = lo.x, qr[g][1] = lo.y, qr[g][2] = lo.z, qr[g][3] = lo.w;
		qr[g][4] = hi.x, qr[g][5] = hi.y, qr[g][6] = hi.z, qr[g][7] = hi.w;
	}
	if (threadIdx.x < G) m_run[threadIdx.x] = -INFINITY, l_run[threadIdx.x] = 0.f;
	float acc[G][8];
#pragma unroll
	for (int g = 0; g < G; ++g)
#pragma unroll
		for (int j = 0; j < 8; ++j) acc[g][j] = 0.f;
	const float inv_div = sqrtf((float)HD);
	const __half *kbase = a.kb + (size_t)kvh * a.max_seq_len * HD + sub * 8;
	const __half *vbase = a.vb + (size_t)kvh * a.max_seq_len * HD + sub * 8;
	__syncthreads();

	for (int tile0 = t_begin; tile0 < t_end; tile0 += kTile) {
		const int tile_n = min(kTile, t_end - tile0);
		// ---- phase 1: scores s[t][g] = q_g . K[t] / sqrt(HD) --------------------------------------
		// loop bounds are warp-uniform (the shuffles below need the whole warp); rows are guarded
		for (int rb = wid * RPW; rb < tile_n; rb += RPC * UN) {
			const int r0 = rb + rgrp;
			uint4 kv[UN];
#pragma unroll
			for (int u = 0; u < UN; ++u) {
				const int r = r0 + u * RPC;
				kv[u] = r < tile_n ? *reinterpret_cast<const uint4 *>(kbase + (size_t)(tile0 + r) * kv_dim)
				                   : make_uint4(0, 0, 0, 0);
			}
#pragma unroll
			for (int u = 0; u < UN; ++u) {
				const int r = r0 + u * RPC;
				float kf[8];
				unpack8(kv[u], kf);
				float s[G];
#pragma unroll
				for (int g = 0; g < G; ++g) {
					float d = 0.f;
#pragma unroll
					for (int j = 0; j < 8; ++j) d = fmaf(qr[g][j], kf[j], d);
					s[g] = d;
				}
#pragma unroll
				for (int o = L / 2; o > 0; o >>= 1)
#pragma unroll
					for (int g = 0; g < G; ++g) s[g] += __shfl_xor_sync(kFull, s[g], o);
				if (sub == 0 && r < tile_n) {
#pragma unroll
					for (int g = 0; g < G; ++g) p_s[r * G + g] = s[g] / inv_div;
				}
			}
		}
		__syncthreads();
		// ---- online softmax bookkeeping for this tile ----------------------------------------------
		for (int g = 0; g < G; ++g) {
			float mx = -INFINITY;
			for (int r = threadIdx.x; r < tile_n; r += kAttnThreads) mx = fmaxf(mx, p_s[r * G + g]);
			mx = block_max(mx, red);
			const float m_old = m_run[g];
			const float m_new = fmaxf(m_old, mx);
			float sum = 0.f;
			for (int r = threadIdx.x; r < tile_n; r += kAttnThreads) {
				float e = expf(p_s[r * G + g] - m_new);
				p_s[r * G + g] = e;
				sum += e;
			}
			sum = block_sum(sum, red);
			if (threadIdx.x == 0) {
				const float c = m_old == -INFINITY ? 0.f : expf(m_old - m_new);
				corr_s[g] = c;
				l_run[g] = l_run[g] * c + sum;
				m_run[g] = m_new;
			}
		}
		__syncthreads();
#pragma unroll
		for (int g = 0; g < G; ++g) {
			const float c = corr_s[g];
#pragma unroll
			for (int j = 0; j < 8; ++j) acc[g][j] *= c;
		}
		// ---- phase 2: acc[g] += p[t][g] * V[t] ------------------------------------------------------
		for (int rb = wid * RPW; rb < tile_n; rb += RPC * UN) {
			const int r0 = rb + rgrp;
			uint4 vv[UN];
#pragma unroll
			for (int u = 0; u < UN; ++u) {
				const int r = r0 + u * RPC;
				vv[u] = r < tile_n ? *reinterpret_cast<const uint4 *>(vbase + (size_t)(tile0 + r) * kv_dim)
				                   : make_uint4(0, 0, 0, 0);
			}
#pragma unroll
			for (int u = 0; u < UN; ++u) {
				const int r = r0 + u * RPC;
				if (r < tile_n) {
					float vf[8];
					unpack8(vv[u], vf);
#pragma unroll
					for (int g = 0; g < G; ++g) {
						const float p = p_s[r * G + g];
#pragma unroll
						for (int j = 0; j < 8; ++j) acc[g][j] = fmaf(p, vf[j], acc[g][j]);
					}
				}
			}
		}
		__syncthreads(); // p_s is rewritten by the next tile
	}

	// ---- reduce the CTA's accumulators: row groups inside a warp, then warps -----------------------
#pragma unroll
	for (int o = 16; o >= L; o >>= 1)
#pragma unroll
		for (int g = 0; g < G; ++g)
#pragma unroll
			for (int j = 0; j < 8; ++j) acc[g][j] += __shfl_xor_sync(kFull, acc[g][j], o);
	if (rgrp == 0) {
#pragma unroll
		for (int g = 0; g < G; ++g)
#pragma unroll
			for (int j = 0; j < 8; ++j) o_s[(wid * G + g) * HD + sub * 8 + j] = acc[g][j];
	}
	__syncthreads();
	float *my_acc = a.ws_acc + ((size_t)kvh * S + split) * G * HD;
	for (int i = threadIdx.x; i < G * HD; i += kAttnThreads) {
		float v = 0.f;
#pragma unroll
		for (int w = 0; w < kAttnWarps; ++w) v += o_s[w * G * HD + i];
		my_acc[i] = v;
	}
	if (threadIdx.x < G) {
		float *ml = a.ws_ml + (((size_t)kvh * S + split) * G + threadIdx.x) * 2;
		ml[0] = m_run[threadIdx.x];
		ml[1] = l_run[threadIdx.x];
	}
	// ---- last CTA of this KV head merges the splits -------------------------------------------------
	__threadfence();
	__syncthreads();
	if (threadIdx.x == 0) {
		unsigned prev = atomicAdd(&a.counters[kvh], 1u);
		is_last = prev == (unsigned)(S - 1);
	}
	__syncthreads();
	if (!is_last) return;
	__threadfence();
	float *w_s = p_s; // [S][G] merge weights (S <= kTile)
	for (int g = wid; g < G; g += kAttnWarps) {
		const float *ml = a.ws_ml + ((size_t)kvh * S * G + g) * 2;
		float M = -INFINITY;
		for (int s = lane; s < S; s += 32) M = fmaxf(M, __ldcg(ml + (size_t)s * G * 2));
		M = warp_max(M);
		float Lsum = 0.f;
		for (int s = lane; s < S; s += 32) {
			const float m = __ldcg(ml + (size_t)s * G * 2), l = __ldcg(ml + (size_t)s * G * 2 + 1);
			const float w = m == -INFINITY ? 0.f : expf(m - M);
			w_s[s * G + g] = w;
			Lsum += w * l;
		}
		Lsum = warp_sum(Lsum);
		if (lane == 0) l_run[g] = Lsum;
	}
	__syncthreads();
	for (int i = threadIdx.x; i < G * HD; i += kAttnThreads) {
		const int g = i / HD;
		const float *pa = a.ws_acc + (size_t)kvh * S * G * HD + i;
		float v = 0.f;
		for (int s = 0; s < S; ++s) v = fmaf(w_s[s * G + g], __ldcg(pa + (size_t)s * G * HD), v);
		a.out[(size_t)kvh * G * HD + i] = v / l_run[g];
	}
	if (threadIdx.x == 0) a.counters[kvh] = 0; // ready for the next launch
}

// Generic path: any head_dim / group size, one CTA per q head, materialises the normalised attention
// row in `att` exactly like the reference (src/infer.cu:338-384) — used for shapes the fast path does
// not cover and by the mha test entry, which must return `att` (src/test.cpp:171-188).
__global__ void __launch_bounds__(256) attn_generic_kernel(float *out, float *att, const __half *kb, const __half *vb,
                                                           const float *q, int head_dim, int n_heads, int n_kv_heads,
                                                           int max_seq_len, int kv_len_arg, const int *ctl) {
	__shared__ float red[36];
	const int h = blockIdx.x, kvh = h / (n_heads / n_kv_heads);
	const int kv_len = ctl ? ctl[CTL_KVLEN] : kv_len_arg;
	const __half *kh = kb + (size_t)kvh * max_seq_len * head_dim, *vh = vb + (size_t)kvh * max_seq_len * head_dim;
	const float *qh = q + (size_t)h * head_dim;
	float *atth = att + (size_t)h * max_seq_len;
	float mx = -INFINITY;
	for (int t = threadIdx.x; t < kv_len; t += blockDim.x) {
		const __half *k = kh + (size_t)t * head_dim;
		float s = 0.f;
		for (int i = 0; i < head_dim; ++i) s += qh[i] * __half2float(k[i]);
		s /= sqrtf((float)head_dim);
		atth[t] = s;
		mx = fmaxf(mx, s);
	}
	mx = block_max(mx, red);
	float sum = 0.f;
	for (int t = threadIdx.x; t < kv_len; t += blockDim.x) {
		float e = expf(atth[t] - mx);
		atth[t] = e;
		sum += e;
	}
	sum = block_sum(sum, red);
	for (int t = threadIdx.x; t < kv_len; t += blockDim.x) atth[t] /= sum;
	__syncthreads();
	for (int i = threadIdx.x; i < head_dim; i += blockDim.x) {
		float v = 0.f;
		for (int t = 0; t < kv_len; ++t) v += atth[t] * __half2float(vh[(size_t)t * head_dim + i]);
		out[(size_t)h * head_dim + i] = v;
	}
}

static int attn_splits(int n_kv_heads) {
	int s = 2 * device_info().sms / std::max(1, n_kv_heads);
	return std::max(1, std::min(s, kTile));
}

static bool attn_fast_shape(int head_dim, int g) {
	return (head_dim == 16 || head_dim == 32 || head_dim == 64 || head_dim == 128) &&
	       (g == 1 || g == 2 || g == 4 || g == 8);
}

template <int HD, int G>
static int launch_attn(const AttnArgs &a, int S, cudaStream_t st) {
	const size_t smem = (36 + (size_t)kTile * G + (size_t)kAttnWarps * G * HD) * sizeof(float);
	static bool configured = false;
	if (!configured && smem > 48 * 1024) {
		YALM_TRY(cudaFuncSetAttribute(attn_decode_kernel<HD, G>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
		configured = true;
	}
	attn_decode_kernel<HD, G><<<dim3(S, a.n_kv_heads), kAttnThreads, smem, st>>>(a);
	return (int)cudaGetLastError();
}

} // namespace yalm

using namespace yalm;

extern "C" size_t yalm_attn_workspace_bytes(int n_heads, int n_kv_heads, int head_dim) {
	if (n_heads <= 0 || n_kv_heads <= 0 || head_dim <= 0) return 0;
	const int S = attn_splits(n_kv_heads);
	const size_t g = n_heads / n_kv_heads;
	size_t floats = (size_t)n_kv_heads * S * g * (head_dim + 2);
	return floats * sizeof(float) + (size_t)n_kv_heads * sizeof(unsigned) + 256;
}

extern "C" int yalm_attn_decode(float *out, const uint16_t *key_cache, const uint16_t *value_cache, const float *q,
                                void *workspace, int n_heads, int n_kv_heads, int head_dim, int max_seq_len,
                                const int *ctl, yalm_stream_t s) {
	if (!out || !key_cache || !value_cache || !q || !workspace || !ctl) return YALM_EINVAL;
	if (n_heads <= 0 || n_kv_heads <= 0 || n_heads % n_kv_heads || max_seq_len <= 0) return YALM_EINVAL;
	const int g = n_heads / n_kv_heads;
	if (!attn_fast_shape(head_dim, g)) return YALM_EINVAL; // callers use yalm_attn_generic for odd shapes
	const int S = attn_splits(n_kv_heads);
	AttnArgs a;
	a.out = out, a.kb = (const __half *)key_cache, a.vb = (const __half *)value_cache, a.q = q;
	// counters first (256-byte slot), then ml, then acc
	a.counters = (unsigned *)workspace;
	a.ws_ml = (float *)((char *)workspace + 256);
	a.ws_acc = a.ws_ml + (size_t)n_kv_heads * S * g * 2;
	a.n_kv_heads = n_kv_heads, a.max_seq_len = max_seq_len, a.ctl = ctl;
	cudaStream_t st = (cudaStream_t)s;
#define YALM_ATTN_CASE(HD, G)                                                                                 \
	if (head_dim == HD && g == G) return launch_attn<HD, G>(a, S, st);
	YALM_ATTN_CASE(128, 1) YALM_ATTN_CASE(128, 2) YALM_ATTN_CASE(128, 4) YALM_ATTN_CASE(128, 8)
	YALM_ATTN_CASE(64, 1) YALM_ATTN_CASE(64, 2) YALM_ATTN_CASE(64, 4) YALM_ATTN_CASE(64, 8)
	YALM_ATTN_CASE(32, 1) YALM_ATTN_CASE(32, 2) YALM_ATTN_CASE(32, 4) YALM_ATTN_CASE(32, 8)
	YALM_ATTN_CASE(16, 1) YALM_ATTN_CASE(16, 2) YALM_ATTN_CASE(16, 4) YALM_ATTN_CASE(16, 8)
#undef YALM_ATTN_CASE
	return YALM_EINVAL;
}

// ---- yalm_attn_decode with no allocation anywhere (no counterpart in the reference; added
// for callers that cannot call cudaMalloc: CUDA graph capture, tight serving loops, harnesses that forbid it).
// The split partials live in ONE file-scope device array per GPU that loads the library (held until exit), so:
//   * shapes whose workspace exceeds kAttnStaticBytes are refused (YALM_EINVAL + a line on stderr);
//   * every call on a device shares the array: calls that overlap on the GPU (two streams, or two host threads) race;
//   * the array keeps the last call's contents; only the arrival counters must start at zero, and they are cleared on the
//     caller's stream before each launch (a memset node under capture) — the partials are fully rewritten before being read.
// In host code `g_attn_scratch` is only a stand-in for the symbol: the device address comes from cudaGetSymbolAddress, once
// per device, with that device current.
constexpr size_t kAttnStaticBytes = 2u << 20; // 2 MiB: covers every fast shape on a <= 148-SM part (see below)
constexpr int kAttnStaticMaxDevices = 64;
static __device__ __align__(256) unsigned char g_attn_scratch[kAttnStaticBytes];

static int attn_static_scratch(void **out) {
	static std::mutex mu;
	static void *ptr[kAttnStaticMaxDevices] = {};
	int dev = -1;
	YALM_TRY(cudaGetDevice(&dev));
	if (dev < 0 || dev >= kAttnStaticMaxDevices) return YALM_EINVAL;
	std::lock_guard<std::mutex> lock(mu);
	if (!ptr[dev]) {
		void *p = nullptr;
		YALM_TRY(cudaGetSymbolAddress(&p, g_attn_scratch)); // device `dev` is current: its copy of the array
		ptr[dev] = p;
	}
	*out = ptr[dev];
	return 0;
}

extern "C" size_t yalm_attn_decode_static_max_bytes(void) {
	return kAttnStaticBytes;
}

extern "C" int yalm_attn_decode_static(float *out, const uint16_t *key_cache, const uint16_t *value_cache, const float *q,
                                       int n_heads, int n_kv_heads, int head_dim, int max_seq_len, const int *ctl,
                                       yalm_stream_t s) {
	if (!out || !key_cache || !value_cache || !q || !ctl) return YALM_EINVAL;
	if (n_heads <= 0 || n_kv_heads <= 0 || n_heads % n_kv_heads || max_seq_len <= 0) return YALM_EINVAL;
	if (!attn_fast_shape(head_dim, n_heads / n_kv_heads)) return YALM_EINVAL;
	const size_t need = yalm_attn_workspace_bytes(n_heads, n_kv_heads, head_dim);
	// the counters share the first 256-byte slot of the workspace: 64 KV heads at most
	if (need > kAttnStaticBytes || (size_t)n_kv_heads * sizeof(unsigned) > 256) {
		fprintf(stderr,
		        "yalm_attn_decode_static: n_heads %d, n_kv_heads %d, head_dim %d needs %zu bytes of split-K scratch (and at most "
		        "64 KV heads); the static scratch holds %zu — use yalm_attn_decode with a caller-owned workspace\n",
		        n_heads, n_kv_heads, head_dim, need, kAttnStaticBytes);
		return YALM_EINVAL;
	}
	void *scratch = nullptr;
	if (int rc = attn_static_scratch(&scratch)) return rc;
	YALM_TRY(cudaMemsetAsync(scratch, 0, 256, (cudaStream_t)s)); // arrival counters (a failed earlier launch may have left them)
	return yalm_attn_decode(out, key_cache, value_cache, q, scratch, n_heads, n_kv_heads, head_dim, max_seq_len, ctl, s);
}

// Not in the public header's hot-path list: generic-shape attention with a materialised `att`.
extern "C" int yalm_attn_generic(float *out, float *att, const uint16_t *key_cache, const uint16_t *value_cache,
                                 const float *q, int n_heads, int n_kv_heads, int head_dim, int max_seq_len,
                                 int kv_len, const int *ctl, yalm_stream_t s) {
	if (!out || !att || !key_cache || !value_cache || !q || n_heads <= 0 || n_kv_heads <= 0 || n_heads % n_kv_heads)
		return YALM_EINVAL;
	attn_generic_kernel<<<n_heads, 256, 0, (cudaStream_t)s>>>(out, att, (const __half *)key_cache,
	                                                          (const __half *)value_cache, q, head_dim, n_heads,
	                                                          n_kv_heads, max_seq_len, kv_len, ctl);
	return (int)cudaGetLastError();
}

extern "C" int yalm_attn_is_fast_shape(int n_heads, int n_kv_heads, int head_dim) {
	return n_kv_heads > 0 && n_heads % n_kv_heads == 0 && attn_fast_shape(head_dim, n_heads / n_kv_heads);
}
