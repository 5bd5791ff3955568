// Whole-token persistent decode kernel for sm_100a (see mega.h for the idea).
//
// Roles inside a CTA (one CTA per SM, cooperative launch so grid barriers are safe):
//   warp 8      producer: walks this CTA's tile list for the WHOLE forward pass and issues
//               cp.async.bulk (1-D TMA) copies global -> shared ring, completion on mbarriers.  It never
//               waits for activations, only for free ring stages, so weight tiles of phase p+1 are in
//               flight while the consumers are still finishing (or barrier-waiting for) phase p.
//   warps 0-7   consumers: per phase  grid barrier -> stage the input vector (rmsnorm fused) into
//               shared memory -> for every tile: wait(full) -> fp32 dot products straight out of shared
//               memory -> release(empty) -> cross-warp reduce -> fused epilogue (clip + RoPE + KV append /
//               residual add / GLU / logits + arg-max).
// Phases per layer: QKV | attention (split-K flash-decode over the head-major KV ring) | wo+residual |
// w1/w3+GLU | w2+residual; then final norm + LM head (+ greedy arg-max and control-block advance).
// Replaces, for one token, everything Model::_forward_cuda_build_graph launches (src/infer.cu:1041-1128,
// 699-888): ~323 kernel nodes -> 1 launch.  Arithmetic (fp32 accumulate, formulas, clip/rope/softmax
// definitions) follows the CPU reference exactly as the per-op kernels do (gemv.cuh, ops_attn.cu).
#pragma once

#include "gemv.cuh"
#include "mega.h"

#include <cfloat>

namespace yalm {
namespace mega {

// A wait longer than this many SM cycles (~4 s) is a bug: trap, never hang.  clock64, not %globaltimer: the
// global timer read costs about a microsecond, which on the producer's empty-stage wait throttled the whole ring.
constexpr long long kTimeoutCycles = 8000000000ll;
constexpr int kCT = kConsumerWarps * 32;
constexpr int kEpiBatch = 32; // row groups per epilogue batch (two batches of partials live in shared memory)

// ---- PTX helpers -------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void *p) {
	return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
	asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
	asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
	asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint32_t bar, uint32_t parity) {
	uint32_t ok;
	asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
	             : "=r"(ok)
	             : "r"(bar), "r"(parity)
	             : "memory");
	return ok != 0;
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
	if (mbar_try_wait(bar, parity)) return;
	const long long t0 = clock64();
	while (!mbar_try_wait(bar, parity))
		if (clock64() - t0 > kTimeoutCycles) __trap();
}
// 1-D bulk copy global -> shared, completion counted in bytes on `bar`; weights/KV are read once per
// token, so they are tagged evict-first in L2 (activations and partials stay resident).
__device__ __forceinline__ void bulk_g2s(uint32_t dst, const void *src, uint32_t bytes, uint32_t bar, uint64_t policy) {
	if (policy == 0) return; // debug: no copies
	asm volatile(
	    "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;" ::"r"(dst),
	    "l"(src), "r"(bytes), "r"(bar), "l"(policy)
	    : "memory");
}
__device__ __forceinline__ uint64_t policy_evict_first() {
	uint64_t p;
	asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(p));
	return p;
}
__device__ __forceinline__ void cbar() { // consumer warps only
	asm volatile("bar.sync 1, %0;" ::"n"(kCT) : "memory");
}
__device__ __forceinline__ unsigned ld_acquire(const unsigned *p) {
	unsigned v;
	asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
	return v;
}

// ---- tile shapes ------------------------------------------------------------------------------------------
enum Kind { kQKV = 0, kWO = 1, kFFN = 2, kW2 = 3, kLM = 4, kNumKinds = 5 };

struct Shape {
	int R;  // rows per tile (FFN: R/2 rows of w1 then R/2 rows of w3)
	int KT; // columns per tile
	int nk; // K splits (tiles per row group)
	int G;  // row groups
};

// Rows-per-tile / K-split choice for a GEMV with `rows` output rows (per matrix) over K inputs of B bytes;
// `pair`: rows are consumed two at a time (RoPE pairs; w1/w3 pairs).  Returns false if no shape fits a stage.
__host__ __device__ inline bool tile_shape(int rows, int rows_b, int K, int B, int EP, bool ffn, bool pair, Shape *s) {
	const int CH = 32 * EP;
	if (K <= 0 || rows <= 0 || K % EP) return false;
	for (int R = 4; R >= 1; R >>= 1) {
		if (pair && R < 2) break;
		const int unit = ffn ? R / 2 : R; // rows of one matrix per tile
		if (rows % unit || (rows_b && rows_b % unit)) continue;
		for (int nk = 1; nk <= 64; ++nk) {
			if (K % nk) continue;
			const int KT = K / nk;
			if (nk > 1 && KT % CH) continue;
			if ((long)R * KT * B > kStageBytes) continue;
			s->R = R, s->KT = KT, s->nk = nk;
			return true;
		}
	}
	return false;
}

// Optional phase profile (Args.prof != null): per CTA, per (layer, phase) kProfSlots clock64 values written by
// consumer thread 0: [0] phase entered (before the grid barrier) [1] barrier passed [2] input vector staged
// [3] tiles + epilogue done; [4] cycles warp 0 spent waiting for full ring stages (starvation = HBM-bound).
constexpr int kProfSlots = 8, kProfPhases = 6; // phases: qkv, attn, wo, ffn, w2, lm(+argmax) (last layer slot only)

struct Ctx {
	long long *prof; // this CTA's profile rows or null
	long long wait_cycles;
	int debug;
	unsigned char *stages;
	uint32_t stages_u32, full0, empty0;
	float *vec, *part, *red;
	int *misc;
	int cta, ncta, warp, lane, tid, S;
	uint32_t it; // tiles produced / consumed so far by this role
	int rot;     // CTA rotation of the current GEMV phase
	unsigned bar_gen;
	int token, pos, kv_sink, kv_pos, kv_len;
	Shape shape[kNumKinds];
};

// ---- producer ---------------------------------------------------------------------------------------------
struct Producer {
	Ctx &c;
	uint64_t policy;
	__device__ __forceinline__ uint32_t acquire(uint32_t bytes, uint32_t *bar) {
		const uint32_t s = c.it % c.S, par = ((c.it / c.S) & 1u) ^ 1u;
		mbar_wait(c.empty0 + 8 * s, par);
		*bar = c.full0 + 8 * s;
		if (c.debug & 8) mbar_arrive(*bar);
		else mbar_expect_tx(*bar, bytes);
		++c.it;
		return c.stages_u32 + s * kStageBytes;
	}
	// rows [row0, row0+nrows) x columns [kc*KT, (kc+1)*KT) of a (rows, K) matrix -> dst, tile rows packed
	__device__ __forceinline__ void rows_copy(uint32_t dst, uint32_t bar, const char *m, long row0, int nrows, int K, int KT,
	                                          int kc, int nk, int B) {
		const size_t row_bytes = (size_t)K * B;
		if (nk == 1) {
			bulk_g2s(dst, m + row0 * row_bytes, (uint32_t)(nrows * row_bytes), bar, policy);
		} else {
			const uint32_t tb = (uint32_t)KT * B;
			for (int r = 0; r < nrows; ++r)
				bulk_g2s(dst + r * tb, m + (row0 + r) * row_bytes + (size_t)kc * tb, tb, bar, policy);
		}
	}
	__device__ void gemv(int kind, const char *m0, const char *m1, const char *m2, int rows0, int rows1, int K, int B) {
		const Shape sh = c.shape[kind];
		const uint32_t tile_bytes = (uint32_t)sh.R * sh.KT * B;
		for (int g = (c.cta - c.rot + c.ncta) % c.ncta; g < sh.G; g += c.ncta) {
			for (int kc = 0; kc < sh.nk; ++kc) {
				uint32_t bar;
				const uint32_t dst = acquire(tile_bytes, &bar);
				if (kind == kFFN) {
					const int rh = sh.R / 2;
					rows_copy(dst, bar, m0, (long)g * rh, rh, K, sh.KT, kc, sh.nk, B);
					rows_copy(dst + rh * sh.KT * B, bar, m1, (long)g * rh, rh, K, sh.KT, kc, sh.nk, B);
				} else {
					long row = (long)g * sh.R;
					const char *m = m0;
					if (kind == kQKV) {
						if (row >= rows0 + rows1) row -= rows0 + rows1, m = m2;
						else if (row >= rows0) row -= rows0, m = m1;
					}
					rows_copy(dst, bar, m, row, sh.R, K, sh.KT, kc, sh.nk, B);
				}
			}
		}
		c.rot = (c.rot + sh.G) % c.ncta;
	}
	__device__ void attention(const Layer &L, int HD, int n_kv, int max_seq) {
		const int C = kStageBytes / (4 * HD);
		const int nch = (c.kv_len + C - 1) / C, U = n_kv * nch;
		const int u0 = (int)((long long)c.cta * U / c.ncta), u1 = (int)((long long)(c.cta + 1) * U / c.ncta);
		for (int u = u0; u < u1; ++u) {
			const int kvh = u / nch, t0 = (u % nch) * C, n = min(C, c.kv_len - t0);
			const uint32_t bytes = (uint32_t)n * HD * 2;
			uint32_t bar;
			const uint32_t dst = acquire(2 * bytes, &bar);
			const size_t off = ((size_t)kvh * max_seq + t0) * HD;
			bulk_g2s(dst, L.kc + off, bytes, bar, policy);
			bulk_g2s(dst + kStageBytes / 2, L.vc + off, bytes, bar, policy);
		}
	}
};

// ---- consumer helpers -------------------------------------------------------------------------------------
__device__ __forceinline__ float cons_sum(Ctx &c, float v) {
	v = warp_sum(v);
	cbar();
	if (c.lane == 0) c.red[c.warp] = v;
	cbar();
	float t = 0.f;
#pragma unroll
	for (int w = 0; w < kConsumerWarps; ++w) t += c.red[w];
	return t;
}

__device__ __forceinline__ void grid_barrier(Ctx &c, unsigned *sync) {
	++c.bar_gen;
	cbar();
	if (c.tid == 0) {
		const unsigned target = c.bar_gen * (unsigned)c.ncta;
		__threadfence();
		atomicAdd(sync, 1u);
		if (ld_acquire(sync) < target) {
			const long long t0 = clock64();
			while (ld_acquire(sync) < target)
				if (clock64() - t0 > kTimeoutCycles) __trap();
		}
		__threadfence();
	}
	cbar();
}

__device__ __forceinline__ void stamp(Ctx &c, int layer, int phase, int k) {
	if (c.prof && c.tid == 0) {
		long long *row = c.prof + ((size_t)layer * kProfPhases + phase) * kProfSlots;
		row[k] = clock64();
		if (k == 3) row[4] = c.wait_cycles, c.wait_cycles = 0;
	}
}

__device__ __forceinline__ const unsigned char *wait_full(Ctx &c, uint32_t *slot) {
	const uint32_t s = c.it % c.S, par = (c.it / c.S) & 1u;
	if (c.prof && c.tid == 0) {
		const long long t0 = clock64();
		mbar_wait(c.full0 + 8 * s, par);
		c.wait_cycles += clock64() - t0;
	} else
		mbar_wait(c.full0 + 8 * s, par);
	*slot = s;
	++c.it;
	return c.stages + (size_t)s * kStageBytes;
}
__device__ __forceinline__ void release(Ctx &c, uint32_t slot) {
	__syncwarp();
	if (c.lane == 0) mbar_arrive(c.empty0 + 8 * slot);
}

// Input vector -> permuted shared layout (xs_slot).  src_kind 0: fp32 vector in global memory (read through L2,
// it was written by other CTAs in this launch); 1: embedding row `token` of the weight-typed table (layer 0).
template <int DT>
__device__ __forceinline__ float4 load_in4(const Args &a, const float *src, int embed_token, int e) {
	if (embed_token < 0) return __ldcg(reinterpret_cast<const float4 *>(src + e));
	const size_t base = (size_t)embed_token * a.dim + e;
	const float sc = (DT == YALM_F8E4M3 && a.embed_scale) ? a.embed_scale[embed_token] : 1.0f;
	return make_float4(WTraits<DT>::to_float(a.embed, base) * sc, WTraits<DT>::to_float(a.embed, base + 1) * sc,
	                   WTraits<DT>::to_float(a.embed, base + 2) * sc, WTraits<DT>::to_float(a.embed, base + 3) * sc);
}

template <int DT>
__device__ __forceinline__ void stage_vec(Ctx &c, const Args &a, const float *src, int n, const float *rms_w, int embed_token) {
	constexpr int EP = WTraits<DT>::EP;
	constexpr int NC = 8; // float4 per thread kept in registers between the two rmsnorm passes (n <= 8192)
	if (rms_w) {
		// reference formula and association order (src/infer.cpp:134-144): o = x * (1/sqrt(ss/n + eps)) * w
		float4 cache[NC], wv[NC];
		float ss = 0.f;
#pragma unroll
		for (int i = 0; i < NC; ++i) {
			const int e = (c.tid + i * kCT) * 4;
			if (e < n) {
				cache[i] = load_in4<DT>(a, src, embed_token, e);
				wv[i] = *reinterpret_cast<const float4 *>(rms_w + e);
			}
		}
#pragma unroll
		for (int i = 0; i < NC; ++i) {
			const int e = (c.tid + i * kCT) * 4;
			if (e < n) ss += cache[i].x * cache[i].x + cache[i].y * cache[i].y + cache[i].z * cache[i].z + cache[i].w * cache[i].w;
		}
		for (int e = (c.tid + NC * kCT) * 4; e < n; e += kCT * 4) {
			const float4 v = load_in4<DT>(a, src, embed_token, e);
			ss += v.x * v.x + v.y * v.y + v.z * v.z + v.w * v.w;
		}
		ss = cons_sum(c, ss);
		const float scale = 1.0f / sqrtf(ss / n + a.eps);
		const bool seed_x = embed_token >= 0 && c.cta == 0; // layer 0: the residual stream starts as the embedding row
#pragma unroll
		for (int i = 0; i < NC; ++i) {
			const int e = (c.tid + i * kCT) * 4;
			if (e < n) {
				const float4 v = cache[i], w = wv[i];
				*reinterpret_cast<float4 *>(c.vec + xs_slot<EP>(e)) =
				    make_float4(v.x * scale * w.x, v.y * scale * w.y, v.z * scale * w.z, v.w * scale * w.w);
				if (seed_x) *reinterpret_cast<float4 *>(a.x + e) = v;
			}
		}
		for (int e = (c.tid + NC * kCT) * 4; e < n; e += kCT * 4) {
			const float4 v = load_in4<DT>(a, src, embed_token, e);
			const float4 w = *reinterpret_cast<const float4 *>(rms_w + e);
			*reinterpret_cast<float4 *>(c.vec + xs_slot<EP>(e)) =
			    make_float4(v.x * scale * w.x, v.y * scale * w.y, v.z * scale * w.z, v.w * scale * w.w);
			if (seed_x) *reinterpret_cast<float4 *>(a.x + e) = v;
		}
	} else {
		int e = c.tid * 4;
		for (; e + 3 * kCT * 4 < n; e += 4 * kCT * 4) { // four independent L2 loads in flight per thread
			float4 v[4];
#pragma unroll
			for (int i = 0; i < 4; ++i) v[i] = __ldcg(reinterpret_cast<const float4 *>(src + e + i * kCT * 4));
#pragma unroll
			for (int i = 0; i < 4; ++i) *reinterpret_cast<float4 *>(c.vec + xs_slot<EP>(e + i * kCT * 4)) = v[i];
		}
		for (; e < n; e += kCT * 4)
			*reinterpret_cast<float4 *>(c.vec + xs_slot<EP>(e)) = __ldcg(reinterpret_cast<const float4 *>(src + e));
	}
	cbar();
}

// Per-phase epilogue state.
struct Epi {
	int kind;
	const Layer *L;
	const float *scale0, *scale1, *scale2; // fp8 row scales per matrix
	int rows0, rows1;                      // QKV boundaries
	float best_v;                          // LM head: running arg-max of this thread's rows
	int best_i;
};

// Epilogue of output item `t` of row group `g` (a row, a RoPE pair or a w1/w3 pair); `part` holds the eight
// warps' partial sums of the group's rows: part[warp*4 + r].  Sum order over warps is fixed (deterministic).
template <int DT>
__device__ __forceinline__ void finalize(Ctx &c, const Args &a, Epi &ep, const Shape &sh, int g, int t, const float *part) {
	auto rowsum = [&](int r) {
		float v = 0.f;
#pragma unroll
		for (int w = 0; w < kConsumerWarps; ++w) v += part[w * 4 + r];
		return v;
	};
	switch (ep.kind) {
	case kQKV: {
		long row = (long)g * sh.R + 2 * t;
		int which = 0;
		const float *sc = ep.scale0;
		if (row >= ep.rows0 + ep.rows1) which = 2, row -= ep.rows0 + ep.rows1, sc = ep.scale2;
		else if (row >= ep.rows0) which = 1, row -= ep.rows0, sc = ep.scale1;
		float v0 = rowsum(2 * t), v1 = rowsum(2 * t + 1);
		if (DT == YALM_F8E4M3 && sc) v0 *= sc[row], v1 *= sc[row + 1];
		v0 = clipf(v0, a.clip), v1 = clipf(v1, a.clip);
		const int hd = a.head_dim, j = (int)(row % hd);
		if (which < 2) rope_pair(v0, v1, j, c.pos, a.theta, a.rotary_dim);
		if (which == 0) {
			*reinterpret_cast<float2 *>(a.q + row) = make_float2(v0, v1);
		} else {
			__half *cache = reinterpret_cast<__half *>(which == 1 ? ep.L->kc : ep.L->vc);
			__half *dst = cache + ((size_t)(row / hd) * a.max_seq_len + c.kv_pos) * hd + j;
			*reinterpret_cast<__half2 *>(dst) = __floats2half2_rn(v0, v1);
		}
		return;
	}
	case kWO:
	case kW2: {
		const long row = (long)g * sh.R + t;
		float v = rowsum(t);
		if (DT == YALM_F8E4M3 && ep.scale0) v *= ep.scale0[row];
		a.x[row] = __ldcg(a.x + row) + v; // x += W v (src/infer.cpp:335-337, 381-383)
		return;
	}
	case kFFN: {
		const int rh = sh.R / 2;
		const long row = (long)g * rh + t;
		float gt = rowsum(t), up = rowsum(rh + t);
		if (DT == YALM_F8E4M3 && ep.scale0) gt *= ep.scale0[row], up *= ep.scale1[row];
		gt = a.act == YALM_ACT_GELU ? act_gelu(gt) : act_silu(gt);
		a.hb[row] = gt * up;
		return;
	}
	default: { // kLM
		const long row = (long)g * sh.R + t;
		float v = rowsum(t);
		if (DT == YALM_F8E4M3 && ep.scale0) v *= ep.scale0[row];
		a.logits[row] = v;
		if (v > ep.best_v) ep.best_v = v, ep.best_i = (int)row; // a thread's rows ascend: first maximum wins
		return;
	}
	}
}

// All row groups of one GEMV phase owned by this CTA.  The input vector is already staged in c.vec.
// The eight warps never wait for each other inside the tile loop: each adds its slice of every tile and parks
// its per-row partial in shared memory; every kEpiBatch groups (and at the end) one barrier hands a whole batch
// to all 256 threads for the cross-warp sum and the fused epilogue, so global-memory latency in the epilogue
// (residual read, KV append) is paid once per batch, not once per tile.
template <int DT, int R>
__device__ __forceinline__ void gemv_consume(Ctx &c, const Args &a, Epi &ep) {
	using T = WTraits<DT>;
	constexpr int EP = T::EP, CH = 32 * EP, Q = EP / 4;
	const Shape sh = c.shape[ep.kind];
	const int chunks = (sh.KT + CH - 1) / CH;
	const int row_stride = sh.KT * T::BYTES;
	const int lane = c.lane, warp = c.warp;
	const int nfin = (ep.kind == kQKV || ep.kind == kFFN) ? sh.R / 2 : sh.R; // epilogue items per group
	const int g_first = (c.cta - c.rot + c.ncta) % c.ncta;
	auto flush = [&](int batch, int n_groups) { // epilogue of groups [batch*kEpiBatch, +n_groups) of this CTA
		const float *pb = c.part + (batch & 1) * kEpiBatch * kConsumerWarps * 4;
		for (int i = c.tid; i < n_groups * nfin; i += kCT) {
			const int gl = i / nfin, t = i % nfin;
			finalize<DT>(c, a, ep, sh, g_first + (batch * kEpiBatch + gl) * c.ncta, t, pb + gl * kConsumerWarps * 4);
		}
	};
	int gi = 0; // local group index
	for (int g = g_first; g < sh.G; g += c.ncta, ++gi) {
		float acc[R];
#pragma unroll
		for (int r = 0; r < R; ++r) acc[r] = 0.f;
		for (int kc = 0; kc < sh.nk; ++kc) {
			uint32_t slot;
			const unsigned char *tile = wait_full(c, &slot);
			const int cbase = kc * chunks;
			// chunk (cbase + cc) belongs to warp (cbase + cc) % 8: balanced over the K splits
			int cc = (warp - cbase % kConsumerWarps + kConsumerWarps) % kConsumerWarps;
			if (c.debug & 1) cc = chunks;
#pragma unroll 2
			for (; cc < chunks; cc += kConsumerWarps) {
				if (cc * CH + lane * EP < sh.KT) {
					uint4 w[R];
#pragma unroll
					for (int r = 0; r < R; ++r)
						w[r] = *reinterpret_cast<const uint4 *>(tile + r * row_stride + cc * 512 + lane * 16);
					float4 xv[Q];
#pragma unroll
					for (int q = 0; q < Q; ++q)
						xv[q] = *reinterpret_cast<const float4 *>(c.vec + (size_t)(cbase + cc) * CH + q * 128 + lane * 4);
#pragma unroll
					for (int r = 0; r < R; ++r) acc[r] = T::fma16(w[r], xv, acc[r]);
				}
			}
			release(c, slot);
		}
#pragma unroll
		for (int r = 0; r < R; ++r) acc[r] = warp_sum(acc[r]);
		if (lane == 0) {
			float *dst = c.part + ((((gi / kEpiBatch) & 1) * kEpiBatch + gi % kEpiBatch) * kConsumerWarps + warp) * 4;
#pragma unroll
			for (int r = 0; r < R; ++r) dst[r] = acc[r];
		}
		if ((gi + 1) % kEpiBatch == 0) {
			cbar(); // batch complete; also fences the flush two batches ago off the buffer half about to be refilled
			flush(gi / kEpiBatch, kEpiBatch);
		}
	}
	cbar();
	if (gi % kEpiBatch) flush(gi / kEpiBatch, gi % kEpiBatch);
	c.rot = (c.rot + sh.G) % c.ncta;
}

template <int DT>
__device__ __forceinline__ void gemv_phase(Ctx &c, const Args &a, Epi &ep) {
	const int R = c.shape[ep.kind].R;
	if (R == 4) gemv_consume<DT, 4>(c, a, ep);
	else if (R == 2) gemv_consume<DT, 2>(c, a, ep);
	else gemv_consume<DT, 1>(c, a, ep);
}

// ---- attention --------------------------------------------------------------------------------------------
__device__ __forceinline__ void unpack8(uint4 v, float (&f)[8]) {
	float2 a = __half22float2(*reinterpret_cast<const __half2 *>(&v.x));
	float2 b = __half22float2(*reinterpret_cast<const __half2 *>(&v.y));
	float2 c = __half22float2(*reinterpret_cast<const __half2 *>(&v.z));
	float2 d = __half22float2(*reinterpret_cast<const __half2 *>(&v.w));
	f[0] = a.x, f[1] = a.y, f[2] = b.x, f[3] = b.y, f[4] = c.x, f[5] = c.y, f[6] = d.x, f[7] = d.y;
}

__device__ __forceinline__ bool unit_overlap(int b, int ncta, int U, int X, int Y) {
	const int b0 = (int)((long long)b * U / ncta), b1 = (int)((long long)(b + 1) * U / ncta);
	return max(b0, X) < min(b1, Y);
}

// Split-K flash-decode over this CTA's (kv head, row chunk) units; arithmetic as src/infer.cpp:216-248.
// Each warp keeps its own online-softmax state over the rows it sees; warps, then CTAs, are merged with the
// usual (max, sum, acc) rescaling in a fixed order, so results are run-to-run identical.
template <int HD, int G>
__device__ __forceinline__ void attention_consume(Ctx &c, const Args &a, const Layer &L) {
	constexpr int LPR = HD / 8, RPS = 32 / LPR;
	constexpr int C = kStageBytes / (4 * HD);
	constexpr int STEPS = C / (kConsumerWarps * RPS);
	constexpr int WS = HD + 2; // per (warp, head) record in shared memory: acc[HD], m, l
	static_assert(STEPS >= 1, "stage too small for this head_dim");
	const int lane = c.lane, warp = c.warp, ncta = c.ncta;
	const int sub = lane % LPR, rgrp = lane / LPR;
	const int nch = (c.kv_len + C - 1) / C, U = a.n_kv_heads * nch;
	const int u0 = (int)((long long)c.cta * U / ncta), u1 = (int)((long long)(c.cta + 1) * U / ncta);
	const float div = sqrtf((float)HD);
	float qr[G][8], acc[G][8], m[G], l[G];
	int cur = -1;
	for (int u = u0; u < u1; ++u) {
		const int kvh = u / nch, t0 = (u % nch) * C, n = min(C, c.kv_len - t0);
		if (kvh != cur) {
			cur = kvh;
#pragma unroll
			for (int g = 0; g < G; ++g) {
				const float *qp = a.q + (size_t)(kvh * G + g) * HD + sub * 8;
				const float4 lo = __ldcg(reinterpret_cast<const float4 *>(qp)), hi = __ldcg(reinterpret_cast<const float4 *>(qp + 4));
				qr[g][0] = lo.x, qr[g][1] = lo.y, qr[g][2] = lo.z, qr[g][3] = lo.w;
				qr[g][4] = hi.x, qr[g][5] = hi.y, qr[g][6] = hi.z, qr[g][7] = hi.w;
				m[g] = -INFINITY, l[g] = 0.f;
#pragma unroll
				for (int j = 0; j < 8; ++j) acc[g][j] = 0.f;
			}
		}
		uint32_t slot;
		const unsigned char *tile = wait_full(c, &slot);
		const __half *ks = reinterpret_cast<const __half *>(tile);
		const __half *vs = reinterpret_cast<const __half *>(tile + kStageBytes / 2);
		const size_t goff = ((size_t)kvh * a.max_seq_len + t0) * HD;
		const __half *kg = reinterpret_cast<const __half *>(L.kc) + goff;
		const __half *vg = reinterpret_cast<const __half *>(L.vc) + goff;
		float s[STEPS][G];
#pragma unroll
		for (int st = 0; st < STEPS; ++st) {
			const int row = st * (kConsumerWarps * RPS) + warp * RPS + rgrp, t = t0 + row;
			const bool valid = row < n && !(c.debug & 4);
			// the row appended by this token and the re-rotated sink rows were written after the producer
			// may have fetched the tile: read those from global memory (L2), everything else from the ring
			const bool fresh = t == c.kv_pos || t < c.kv_sink;
			uint4 kv = make_uint4(0, 0, 0, 0);
			if (valid) kv = fresh ? __ldcg(reinterpret_cast<const uint4 *>(kg + (size_t)row * HD + sub * 8))
			                      : *reinterpret_cast<const uint4 *>(ks + row * HD + sub * 8);
			float kf[8];
			unpack8(kv, kf);
#pragma unroll
			for (int g = 0; g < G; ++g) {
				float d = 0.f;
#pragma unroll
				for (int j = 0; j < 8; ++j) d = fmaf(qr[g][j], kf[j], d);
#pragma unroll
				for (int o = LPR / 2; o > 0; o >>= 1) d += __shfl_xor_sync(kFull, d, o);
				s[st][g] = valid ? d / div : -INFINITY;
			}
		}
#pragma unroll
		for (int g = 0; g < G; ++g) {
			float mt = s[0][g];
#pragma unroll
			for (int st = 1; st < STEPS; ++st) mt = fmaxf(mt, s[st][g]);
#pragma unroll
			for (int o = LPR; o < 32; o <<= 1) mt = fmaxf(mt, __shfl_xor_sync(kFull, mt, o));
			const float m_new = fmaxf(m[g], mt);
			const float corr = m_new == -INFINITY ? 1.f : expf(m[g] - m_new);
			m[g] = m_new;
			l[g] *= corr;
#pragma unroll
			for (int j = 0; j < 8; ++j) acc[g][j] *= corr;
#pragma unroll
			for (int st = 0; st < STEPS; ++st) {
				const float p = s[st][g] == -INFINITY ? 0.f : expf(s[st][g] - m_new);
				s[st][g] = p;
				l[g] += p;
			}
		}
#pragma unroll
		for (int st = 0; st < STEPS; ++st) {
			const int row = st * (kConsumerWarps * RPS) + warp * RPS + rgrp, t = t0 + row;
			if (row < n && !(c.debug & 4)) {
				const bool fresh = t == c.kv_pos || t < c.kv_sink;
				const uint4 vv = fresh ? __ldcg(reinterpret_cast<const uint4 *>(vg + (size_t)row * HD + sub * 8))
				                       : *reinterpret_cast<const uint4 *>(vs + row * HD + sub * 8);
				float vf[8];
				unpack8(vv, vf);
#pragma unroll
				for (int g = 0; g < G; ++g)
#pragma unroll
					for (int j = 0; j < 8; ++j) acc[g][j] = fmaf(s[st][g], vf[j], acc[g][j]);
			}
		}
		release(c, slot);
		if (u + 1 < u1 && (u + 1) / nch == kvh) continue;

		// ---- this CTA is done with head kvh: merge row groups, warps, then publish the partial ----------
#pragma unroll
		for (int o = LPR; o < 32; o <<= 1)
#pragma unroll
			for (int g = 0; g < G; ++g) {
				l[g] += __shfl_xor_sync(kFull, l[g], o);
#pragma unroll
				for (int j = 0; j < 8; ++j) acc[g][j] += __shfl_xor_sync(kFull, acc[g][j], o);
			}
		float *wsm = c.vec;
		if (rgrp == 0) {
#pragma unroll
			for (int g = 0; g < G; ++g) {
				float *rec = wsm + (warp * G + g) * WS;
#pragma unroll
				for (int j = 0; j < 8; ++j) rec[sub * 8 + j] = acc[g][j];
				if (sub == 0) rec[HD] = m[g], rec[HD + 1] = l[g];
			}
		}
		cbar();
		float *pa = a.ws_acc + ((size_t)kvh * ncta + c.cta) * G * HD;
		float *pml = a.ws_ml + ((size_t)kvh * ncta + c.cta) * G * 2;
		for (int i = c.tid; i < G * HD; i += kCT) {
			const int g = i / HD, d = i % HD;
			float M = -INFINITY;
#pragma unroll
			for (int w = 0; w < kConsumerWarps; ++w) M = fmaxf(M, wsm[(w * G + g) * WS + HD]);
			float v = 0.f, ls = 0.f;
#pragma unroll
			for (int w = 0; w < kConsumerWarps; ++w) {
				const float *rec = wsm + (w * G + g) * WS;
				const float e = rec[HD] == -INFINITY ? 0.f : expf(rec[HD] - M);
				v = fmaf(e, rec[d], v);
				ls = fmaf(e, rec[HD + 1], ls);
			}
			pa[i] = v;
			if (d == 0) pml[g * 2] = M, pml[g * 2 + 1] = ls;
		}
		__threadfence();
		cbar();
		const int X = kvh * nch, Y = X + nch;
		if (warp == 0) {
			int cnt = 0;
			for (int b = lane; b < ncta; b += 32) cnt += unit_overlap(b, ncta, U, X, Y) ? 1 : 0;
#pragma unroll
			for (int o = 16; o > 0; o >>= 1) cnt += __shfl_xor_sync(kFull, cnt, o);
			if (lane == 0) {
				const unsigned prev = atomicAdd(&a.head_counters[kvh], 1u);
				c.misc[0] = prev + 1 == (unsigned)cnt;
			}
		}
		cbar();
		if (c.misc[0]) {
			// last CTA of this head: combine every CTA's partial in CTA order
			__threadfence();
			const int b_lo = max(0, (int)((long long)X * ncta / U) - 1);
			const int b_hi = min(ncta - 1, (int)(((long long)Y * ncta + U - 1) / U));
			float *wts = c.vec; // [ncta][G] merge weights (-1: CTA has no partial), then Lsum[G]
			for (int g = warp; g < G; g += kConsumerWarps) {
				float M = -INFINITY;
				for (int b = b_lo + lane; b <= b_hi; b += 32)
					if (unit_overlap(b, ncta, U, X, Y)) M = fmaxf(M, __ldcg(a.ws_ml + (((size_t)kvh * ncta + b) * G + g) * 2));
				M = warp_max(M);
				float ls = 0.f;
				for (int b = b_lo + lane; b <= b_hi; b += 32) {
					float w = -1.f;
					if (unit_overlap(b, ncta, U, X, Y)) {
						const float *ml = a.ws_ml + (((size_t)kvh * ncta + b) * G + g) * 2;
						const float mb = __ldcg(ml);
						w = mb == -INFINITY ? 0.f : expf(mb - M);
						ls += w * __ldcg(ml + 1);
					}
					wts[b * G + g] = w;
				}
				ls = warp_sum(ls);
				if (lane == 0) wts[ncta * G + g] = ls;
			}
			cbar();
			for (int i = c.tid; i < G * HD; i += kCT) {
				const int g = i / HD;
				float v = 0.f;
				for (int b = b_lo; b <= b_hi; ++b) {
					const float w = wts[b * G + g];
					if (w > 0.f) v = fmaf(w, __ldcg(a.ws_acc + ((size_t)kvh * ncta + b) * G * HD + i), v);
				}
				a.att_out[(size_t)kvh * G * HD + i] = v / wts[ncta * G + g];
			}
			if (c.tid == 0) a.head_counters[kvh] = 0;
		}
		cbar(); // c.vec is reused by the next head / phase
	}
}

__device__ __forceinline__ void derive_ring(int pos, int max_seq_len, int *kv_sink, int *kv_pos, int *kv_len) {
	// src/infer.cu:1081-1083
	*kv_sink = pos >= max_seq_len ? KV_SINKS : 0;
	*kv_pos = *kv_sink + (pos - *kv_sink) % (max_seq_len - *kv_sink);
	*kv_len = pos >= max_seq_len ? max_seq_len : pos + 1;
}

// ---- the kernel -----------------------------------------------------------------------------------------
template <int DT, int HD, int G>
__global__ void __launch_bounds__(kThreads, 1) decode_mega_kernel(const __grid_constant__ Args a) {
	extern __shared__ __align__(128) unsigned char smem_raw[];
	using T = WTraits<DT>;
	constexpr int B = T::BYTES, EP = T::EP;
	Ctx c;
	c.S = a.stages;
	c.stages = smem_raw;
	c.stages_u32 = smem_u32(smem_raw);
	c.vec = reinterpret_cast<float *>(smem_raw + (size_t)c.S * kStageBytes);
	c.part = c.vec + a.vec_floats;
	c.red = c.part + 2 * kEpiBatch * kConsumerWarps * 4;
	c.misc = reinterpret_cast<int *>(c.red + 16);
	uint64_t *bars = reinterpret_cast<uint64_t *>(c.misc + 16);
	c.full0 = smem_u32(bars);
	c.empty0 = c.full0 + 8 * kMaxStages;
	c.cta = blockIdx.x, c.ncta = gridDim.x;
	c.tid = threadIdx.x, c.warp = threadIdx.x >> 5, c.lane = threadIdx.x & 31;
	c.it = 0, c.rot = 0, c.bar_gen = 0;
	c.prof = a.prof ? a.prof + (size_t)blockIdx.x * a.n_layers * kProfPhases * kProfSlots : nullptr;
	c.wait_cycles = 0;
	c.debug = a.debug;

	c.token = a.token, c.pos = a.pos;
	if (a.token < 0) c.token = a.ctl[CTL_TOKEN], c.pos = a.ctl[CTL_POS];
	derive_ring(c.pos, a.max_seq_len, &c.kv_sink, &c.kv_pos, &c.kv_len);

	constexpr int hd = HD;
	tile_shape(a.q_dim, a.kv_dim, a.dim, B, EP, false, true, &c.shape[kQKV]);
	c.shape[kQKV].G = (a.q_dim + 2 * a.kv_dim) / c.shape[kQKV].R;
	tile_shape(a.dim, 0, a.q_dim, B, EP, false, false, &c.shape[kWO]);
	c.shape[kWO].G = a.dim / c.shape[kWO].R;
	tile_shape(a.hidden, 0, a.dim, B, EP, true, true, &c.shape[kFFN]);
	c.shape[kFFN].G = a.hidden / (c.shape[kFFN].R / 2);
	tile_shape(a.dim, 0, a.hidden, B, EP, false, false, &c.shape[kW2]);
	c.shape[kW2].G = a.dim / c.shape[kW2].R;
	tile_shape(a.vocab, 0, a.dim, B, EP, false, false, &c.shape[kLM]);
	c.shape[kLM].G = a.vocab / c.shape[kLM].R;

	if (threadIdx.x == 0) {
		for (int i = 0; i < c.S; ++i) {
			mbar_init(c.full0 + 8 * i, 1);
			mbar_init(c.empty0 + 8 * i, kConsumerWarps);
		}
		asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
	}
	__syncthreads();

	if (c.warp == kConsumerWarps) {
		// ================================ producer ================================
		if (c.lane != 0) return;
		Producer p{c, (a.debug & 8) ? 0ull : policy_evict_first()};
		for (int l = 0; l < a.n_layers; ++l) {
			const Layer &L = a.layers[l];
			p.gemv(kQKV, L.wq, L.wk, L.wv, a.q_dim, a.kv_dim, a.dim, B);
			p.attention(L, hd, a.n_kv_heads, a.max_seq_len);
			p.gemv(kWO, L.wo, nullptr, nullptr, a.dim, 0, a.q_dim, B);
			p.gemv(kFFN, L.w1, L.w3, nullptr, a.hidden, 0, a.dim, B);
			p.gemv(kW2, L.w2, nullptr, nullptr, a.dim, 0, a.hidden, B);
		}
		if (a.mode != kHydrate) p.gemv(kLM, a.wcls, nullptr, nullptr, a.vocab, 0, a.dim, B);
		return;
	}

	// ================================ consumers ================================
	Epi ep;
	for (int l = 0; l < a.n_layers; ++l) {
		const Layer &L = a.layers[l];
		ep.L = &L;
		// ---- rmsnorm -> q,k,v -> clip -> RoPE -> KV append (+ sink re-rotation) --------------------------
		stamp(c, l, 0, 0);
		if (l > 0) grid_barrier(c, a.sync);
		stamp(c, l, 0, 1);
		stage_vec<DT>(c, a, a.x, a.dim, L.rms_att, l == 0 ? c.token : -1);
		stamp(c, l, 0, 2);
		if (c.kv_sink > 0 && c.cta == c.ncta - 1) {
			// rotate each sink key forward by one position, fp16 -> fp32 -> fp16 (src/infer.cpp:303-317)
			const int hp = hd >> 1, npairs = c.kv_sink * a.n_kv_heads * hp;
			for (int p = c.tid; p < npairs; p += kCT) {
				const int r = p / (a.n_kv_heads * hp), h = (p / hp) % a.n_kv_heads, i = (p % hp) * 2;
				__half2 *ptr = reinterpret_cast<__half2 *>(reinterpret_cast<__half *>(L.kc) + ((size_t)h * a.max_seq_len + r) * hd + i);
				float2 v = __half22float2(*ptr);
				rope_pair(v.x, v.y, i, 1, a.theta, a.rotary_dim);
				*ptr = __floats2half2_rn(v.x, v.y);
			}
		}
		ep.kind = kQKV, ep.scale0 = L.sq, ep.scale1 = L.sk, ep.scale2 = L.sv, ep.rows0 = a.q_dim, ep.rows1 = a.kv_dim;
		gemv_phase<DT>(c, a, ep);
		stamp(c, l, 0, 3);
		// ---- attention -----------------------------------------------------------------------------------
		stamp(c, l, 1, 0);
		grid_barrier(c, a.sync);
		stamp(c, l, 1, 1);
		stamp(c, l, 1, 2);
		attention_consume<HD, G>(c, a, L);
		stamp(c, l, 1, 3);
		// ---- wo + residual -------------------------------------------------------------------------------
		stamp(c, l, 2, 0);
		grid_barrier(c, a.sync);
		stamp(c, l, 2, 1);
		stage_vec<DT>(c, a, a.att_out, a.q_dim, nullptr, -1);
		stamp(c, l, 2, 2);
		ep.kind = kWO, ep.scale0 = L.so;
		gemv_phase<DT>(c, a, ep);
		stamp(c, l, 2, 3);
		// ---- rmsnorm -> w1/w3 -> GLU ---------------------------------------------------------------------
		stamp(c, l, 3, 0);
		grid_barrier(c, a.sync);
		stamp(c, l, 3, 1);
		stage_vec<DT>(c, a, a.x, a.dim, L.rms_ffn, -1);
		stamp(c, l, 3, 2);
		ep.kind = kFFN, ep.scale0 = L.s1, ep.scale1 = L.s3;
		gemv_phase<DT>(c, a, ep);
		stamp(c, l, 3, 3);
		// ---- w2 + residual -------------------------------------------------------------------------------
		stamp(c, l, 4, 0);
		grid_barrier(c, a.sync);
		stamp(c, l, 4, 1);
		stage_vec<DT>(c, a, a.hb, a.hidden, nullptr, -1);
		stamp(c, l, 4, 2);
		ep.kind = kW2, ep.scale0 = L.s2;
		gemv_phase<DT>(c, a, ep);
		stamp(c, l, 4, 3);
	}
	if (a.mode != kHydrate) {
		// ---- final rmsnorm -> LM head (+ greedy arg-max, first maximum wins: src/sampler.cpp:27-38) -----
		stamp(c, a.n_layers - 1, 5, 0);
		grid_barrier(c, a.sync);
		stamp(c, a.n_layers - 1, 5, 1);
		stage_vec<DT>(c, a, a.x, a.dim, a.rms_final, -1);
		stamp(c, a.n_layers - 1, 5, 2);
		ep.kind = kLM, ep.scale0 = a.wcls_scale, ep.best_v = -FLT_MAX, ep.best_i = 0x7fffffff;
		gemv_phase<DT>(c, a, ep);
		if (a.mode == kGreedy) {
			{
				float bv = ep.best_v;
				int bi = ep.best_i;
#pragma unroll
				for (int o = 16; o > 0; o >>= 1) {
					const float ov = __shfl_xor_sync(kFull, bv, o);
					const int oi = __shfl_xor_sync(kFull, bi, o);
					if (ov > bv || (ov == bv && oi < bi)) bv = ov, bi = oi;
				}
				if (c.lane == 0) c.red[c.warp] = bv, c.misc[4 + c.warp] = bi;
			}
			cbar();
			if (c.tid == 0) {
				float bv = c.red[0];
				int bi = c.misc[4];
				for (int t = 1; t < kConsumerWarps; ++t)
					if (c.red[t] > bv || (c.red[t] == bv && c.misc[4 + t] < bi)) bv = c.red[t], bi = c.misc[4 + t];
				a.amax_val[c.cta] = bv, a.amax_idx[c.cta] = bi;
			}
			grid_barrier(c, a.sync);
			if (c.cta == 0 && c.warp == 0) {
				float bv = -FLT_MAX;
				int bi = 0x7fffffff;
				for (int b = c.lane; b < c.ncta; b += 32) {
					const float v = __ldcg(a.amax_val + b);
					const int i = __ldcg(a.amax_idx + b);
					if (v > bv || (v == bv && i < bi)) bv = v, bi = i;
				}
#pragma unroll
				for (int o = 16; o > 0; o >>= 1) {
					const float ov = __shfl_xor_sync(kFull, bv, o);
					const int oi = __shfl_xor_sync(kFull, bi, o);
					if (ov > bv || (ov == bv && oi < bi)) bv = ov, bi = oi;
				}
				if (c.lane == 0) {
					if (bi == 0x7fffffff) bi = 0;
					const int step = a.ctl[CTL_STEP];
					if (a.tokens_out) a.tokens_out[step] = bi;
					a.ctl[CTL_STEP] = step + 1;
					a.ctl[CTL_TOKEN] = bi;
					a.ctl[CTL_POS] = c.pos + 1;
					int ks, kp, kl;
					derive_ring(c.pos + 1, a.max_seq_len, &ks, &kp, &kl);
					a.ctl[CTL_SINK] = ks, a.ctl[CTL_KVPOS] = kp, a.ctl[CTL_KVLEN] = kl;
				}
			}
		}
	}
	if (a.mode != kHydrate) stamp(c, a.n_layers - 1, 5, 3);
	// ---- leave the barrier words zeroed for the next launch ------------------------------------------
	cbar();
	if (c.tid == 0) {
		__threadfence();
		const unsigned prev = atomicAdd(a.sync + 1, 1u);
		if (prev + 1 == (unsigned)c.ncta) {
			a.sync[0] = 0;
			a.sync[1] = 0;
			__threadfence();
		}
	}
}

// (head_dim, q heads per kv head) combinations compiled; others use the per-op kernel path.
template <int DT>
inline const void *kernel_for(int head_dim, int g) {
#define YALM_MEGA_CASE(HD, GG)                                                                                \
	if (head_dim == HD && g == GG) return (const void *)decode_mega_kernel<DT, HD, GG>;
	YALM_MEGA_CASE(128, 4) YALM_MEGA_CASE(128, 8) YALM_MEGA_CASE(64, 4) YALM_MEGA_CASE(32, 4) YALM_MEGA_CASE(32, 2)
#undef YALM_MEGA_CASE
	return nullptr;
}

} // namespace mega
} // namespace yalm
