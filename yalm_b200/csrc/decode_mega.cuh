// Whole-token persistent decode kernel for sm_100a (see mega.h for the idea).
//
// Roles inside a CTA (one CTA per SM, cooperative launch so grid barriers and flag polling are safe):
//   last warp   producer (all 32 lanes, one tile per lane and batch): walks this CTA's tile list for the WHOLE forward pass and
//               issues cp.async.bulk (1-D TMA) copies global -> shared ring, completion on mbarriers.  It never
//               waits for activations, only for free ring stages, so weight tiles of phase p+1 are in
//               flight while the consumers are still finishing (or waiting for the input of) phase p.
//   warps 0-10  consumers (kConsumerWarps).  Per phase: wait for the input vector — by polling its flagged words
//               ("flagged hand-over" below; one grid barrier per layer remains, in front of attention) -> stage it
//               (rmsnorm fused) into shared memory -> tile loop.  In the tile loop every warp OWNS whole row groups
//               (all K-split tiles of a group land on the same warp), so there is no intra-CTA synchronisation at all
//               inside a tile loop: wait(full) -> fp32 dot products straight out of shared memory ->
//               release(empty) -> shuffle reduce -> the warp's own lanes run the fused epilogue (clip +
//               RoPE + KV append / residual add / GLU / logits + arg-max) and store the flagged result words.
// Phases per layer: QKV | attention (split-K flash-decode over the head-major KV ring) | wo+residual |
// w1/w3+GLU | w2+residual; then final norm + LM head (+ greedy arg-max and control-block advance).
// Replaces, for one token, everything Model::_forward_cuda_build_graph launches (src/infer.cu:1041-1128,
// 699-888): ~323 kernel nodes -> 1 launch.  Arithmetic (fp32 accumulate, formulas, clip/rope/softmax
// definitions) follows the CPU reference exactly as the per-op kernels do (gemv.cuh, ops_attn.cu).
//
// Kernel variants (template parameter V, one translation unit each, chosen by mega::plan): 0 = as described above; 1 = hb
// travels w1/w3 -> w2 slice by slice through bulk copies ("pipelined hb" below: shared memory for ring stages on wide models);
// 2 = the GEMV phases on the tensor cores ("tensor-core GEMV" below: an experiment, opt-in).
//
// Code size matters here: the per-layer instruction footprint has to stay inside the SM's instruction cache
// (every phase body runs once per layer), so every phase body has exactly ONE call site (the phase loops below).
// They are force-inlined all the same: a __noinline__ callee taking Ctx& forces the whole context struct into
// local memory, and with the shared-memory carve-out at its maximum every such access is an L2 round trip.
#pragma once

#include "gemv.cuh"
#include "mega.h"

#include <cfloat>

namespace yalm {
namespace mega {

// A wait longer than this many SM cycles (~60 s) is a bug: trap, never hang.  Long enough for tensor-parallel launches,
// where every local wait is transitively a wait for the slowest rank (ranks may enter a launch seconds apart).
constexpr long long kTimeoutCycles = 120000000000ll;
// Failure record: the first thread whose wait times out says WHICH wait it was before it traps — into pinned host memory,
// which survives the dead context (mega.h: fault_record / describe_fault; the host prints it when a launch fails).
// [0] code (kFault*), [1] CTA, [2] thread, [3] / [4] two wait-specific values.  One pointer per translation unit.
static __device__ unsigned *g_fault_rec = nullptr;
static __device__ long long g_timeout_cycles = kTimeoutCycles; // YALM_MEGA_TIMEOUT_S shortens it (debugging)
enum { kFaultRing = 1, kFaultProducerRel = 2, kFaultConsumerRel = 3, kFaultPoll = 4, kFaultGridBarrier = 5, kFaultTpBarrier = 6,
       kFaultTpMailbox = 7, kFaultRouter = 8, kFaultSample = 9, kFaultHb = 10 };
static __device__ __noinline__ void die(unsigned code, unsigned x, unsigned y) {
	unsigned *r = g_fault_rec;
	if (r && atomicCAS(r, 0u, code) == 0u) {
		r[1] = blockIdx.x, r[2] = threadIdx.x, r[3] = x, r[4] = y;
		__threadfence_system();
	}
	__trap();
}
__device__ __forceinline__ bool timed_out(long long t0) { return clock64() - t0 > g_timeout_cycles; }
inline void set_fault_pointers(unsigned *rec, long long timeout_cycles) { // host: called once per translation unit
	cudaMemcpyToSymbol(g_fault_rec, &rec, sizeof(rec));
	if (timeout_cycles > 0) cudaMemcpyToSymbol(g_timeout_cycles, &timeout_cycles, sizeof(timeout_cycles));
}
constexpr int kCT = kConsumerWarps * 32;
constexpr int kHeadSlots = 1; // a CTA's attention units belong to one KV head (attn_split; needs n_kv_heads <= grid)

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
		if (timed_out(t0)) die(kFaultRing, bar, parity);
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
// The same without a cache hint (activations: every CTA reads them, they must stay in L2).
__device__ __forceinline__ void bulk_g2s_plain(uint32_t dst, const void *src, uint32_t bytes, uint32_t bar) {
	asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst), "l"(src), "r"(bytes),
	             "r"(bar)
	             : "memory");
}
// 3-D tensor copy global -> shared (tensor-core GEMV: 64 columns x rows x column blocks of a weight matrix, SWIZZLE_128B)
__device__ __forceinline__ void tma_load_3d(uint32_t dst, const void *map, int c0, int c1, int c2, uint32_t bar, uint64_t policy) {
	if (policy == 0) return; // debug: no copies
	asm volatile(
	    "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1, {%2, %3, %4}], [%5], %6;" ::"r"(dst),
	    "l"(map), "r"(c0), "r"(c1), "r"(c2), "r"(bar), "l"(policy)
	    : "memory");
}
__device__ __forceinline__ void ldmatrix_x4(uint32_t (&r)[4], uint32_t addr) {
	asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0, %1, %2, %3}, [%4];" : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(addr));
}
__device__ __forceinline__ uint2 lds64(uint32_t addr) {
	uint2 v;
	asm volatile("ld.shared.v2.u32 {%0,%1}, [%2];" : "=r"(v.x), "=r"(v.y) : "r"(addr));
	return v;
}
// D (16 x 8, fp32) += A (16 x 16 fp16, row) * B (16 x 8 fp16, col)
__device__ __forceinline__ void mma16816(float (&d)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
	asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0, %1, %2, %3}, {%4, %5, %6, %7}, {%8, %9}, {%0, %1, %2, %3};"
	             : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
	             : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
// Fire-and-forget L2 prefetch of a global range (no completion tracking, never blocks the issuing thread).
__device__ __forceinline__ void bulk_prefetch_l2(const void *src, uint32_t bytes, bool tma = false) {
	if (tma) { // one request to the copy engine (debug & 16)
		asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(src), "r"(bytes) : "memory");
		return;
	}
	{
	// per-line prefetches through the load/store unit: the bulk form goes through the copy engine, which then
	// handles every byte twice (measured: slower than no prefetch at all)
	const char *p = static_cast<const char *>(src);
	for (uint32_t o = 0; o < bytes; o += 128) asm volatile("prefetch.global.L2 [%0];" ::"l"(p + o) : "memory");
	}
}
__device__ __forceinline__ uint64_t policy_evict_first() {
	uint64_t p;
	asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(p));
	return p;
}
// Per-stage release counters (shared memory): rel[s] = number of times stage s has been consumed and handed
// back.  Tile idx (round r = idx / S of stage s = idx % S) may be issued once rel[s] >= r, and a consumer only
// starts its parity wait on full[s] once rel[s] >= r, i.e. once every earlier round of the stage has completed —
// an mbarrier parity wait alone cannot tell round r from round r-2, and here neither the producers (several
// warps) nor the consumers (each warp owns different tiles) walk a stage's rounds in program order.
// (Plain volatile accesses: ld.acquire / st.release on shared memory made ptxas emit CCTL.IVALL, an L1 invalidate,
// per tile; the ordering that matters is carried by the mbarrier operations that follow the counter accesses.)
__device__ __forceinline__ int ld_rel(const int *p) {
	int v;
	asm volatile("ld.volatile.shared.s32 %0, [%1];" : "=r"(v) : "r"(smem_u32(p)) : "memory");
	return v;
}
__device__ __forceinline__ void st_rel(int *p, int v) {
	asm volatile("st.volatile.shared.s32 [%0], %1;" ::"r"(smem_u32(p)), "r"(v) : "memory");
}
__device__ __forceinline__ void cbar() { // consumer warps only
	asm volatile("bar.sync 1, %0;" ::"n"(kCT) : "memory");
}
__device__ __forceinline__ unsigned ld_acquire(const unsigned *p) {
	unsigned v;
	asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
	return v;
}

// ---- flagged hand-over (single GPU) -----------------------------------------------------------------------------
// Between two phases of a layer a vector (x, hb, att_out, the attention partials) changes hands from the CTAs that
// computed its rows to every CTA.  Instead of "all stores -> CTA barrier -> release counter -> poll -> fence -> CTA
// barrier -> load", every element travels as ONE 64-bit word {fp32 value, u32 flag} written with a single store and
// the readers poll the words themselves: a word whose flag matches is the value (64-bit accesses are single-copy
// atomic), so the hand-over costs one L2 round trip after the last store lands and needs no fence.  flag =
// ll_flag(launch number, producing phase + 1) = seq << kFlagPhaseBits | phase + 1.  It is never 0 (fresh buffers never
// match) and differs between any two launches less than 2^(32 - kFlagPhaseBits) = 1M apart; every word of a buffer is
// rewritten by every launch, so the only words a poller can meet are this launch's and the previous launch's: the
// launch counter may wrap (tests seed it just below the wrap).  plan() rejects models with >= 2^kFlagPhaseBits phases.  Reuse of a buffer is safe because ONE grid barrier per layer
// remains (QKV -> attention, where the appended K/V rows change hands anyway): no CTA can run a layer ahead of
// another one, x alternates between two buffers (after wo / after w2), and hb, att_out and the partials are written
// once per layer on the far side of that barrier from their previous readers.
#ifndef YALM_MEGA_LL
#define YALM_MEGA_LL 1
#endif
// 1: the attention partials are flagged too and every participating CTA combines a slice of the head's outputs
// (ll_combine); 0: arrival counter + combine on the last CTA of the head (combine_head), only att_out is flagged.
// Every CTA reads every word of a flagged vector at the same moment, and an L2 slice hands out a line to one reader at a
// time: measured 1.5 us for the 32 KB of x and 5 us for the 112 KB of hb, three times the aggregate L2 rate.  So the
// producers write kLLRep copies (different lines, different slices; the copies cost no instructions — the otherwise
// idle lanes of the epilogue store them) and CTA c reads copy c % kLLRep.
#ifndef YALM_MEGA_LL_NS
#define YALM_MEGA_LL_NS 32 // first back-off of a polling thread in ns (doubles up to x16)
#endif
#ifndef YALM_MEGA_LL_FIRST
#define YALM_MEGA_LL_FIRST 0 // 1: a thread polls only its first float4 until it is there, then loads the rest (A/B)
#endif
#ifndef YALM_MEGA_LL_SAMPLE
#define YALM_MEGA_LL_SAMPLE 0 // 1: one warp samples the vector first, the others wait at the CTA barrier (A/B: 333 vs 341 tok/s once the vectors are replicated)
#endif
#ifndef YALM_MEGA_LL_ATTN
#define YALM_MEGA_LL_ATTN 0
#endif
__device__ __forceinline__ unsigned ll_flag(unsigned seq, int phase_plus_1) {
	return seq << kFlagPhaseBits | (unsigned)phase_plus_1;
}
__device__ __forceinline__ unsigned long long ll_pack(float v, unsigned flag) {
	return (unsigned long long)__float_as_uint(v) | ((unsigned long long)flag << 32);
}
__device__ __forceinline__ void ll_store(unsigned long long *p, float v, unsigned flag) {
	asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(ll_pack(v, flag)) : "memory");
}
__device__ __forceinline__ void ll_store2(unsigned long long *p, float v0, float v1, unsigned flag) {
	asm volatile("st.relaxed.gpu.global.v2.u64 [%0], {%1, %2};" ::"l"(p), "l"(ll_pack(v0, flag)), "l"(ll_pack(v1, flag)) : "memory");
}
__device__ __forceinline__ ulonglong2 ll_load2(const unsigned long long *p) {
	ulonglong2 v;
	asm volatile("ld.relaxed.gpu.global.v2.u64 {%0, %1}, [%2];" : "=l"(v.x), "=l"(v.y) : "l"(p) : "memory");
	return v;
}
__device__ __forceinline__ unsigned long long ll_load1(const unsigned long long *p) {
	unsigned long long v;
	asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
	return v;
}
// The same words across GPUs (tensor parallelism): stores into a peer's region over NVLink, polls of this rank's own region.
__device__ __forceinline__ void ll_store_sys(unsigned long long *p, float v, unsigned flag) {
	asm volatile("st.relaxed.sys.global.u64 [%0], %1;" ::"l"(p), "l"(ll_pack(v, flag)) : "memory");
}
__device__ __forceinline__ void ll_store2_sys(unsigned long long *p, float v0, float v1, unsigned flag) {
	asm volatile("st.relaxed.sys.global.v2.u64 [%0], {%1, %2};" ::"l"(p), "l"(ll_pack(v0, flag)), "l"(ll_pack(v1, flag)) : "memory");
}
__device__ __forceinline__ unsigned long long ll_load1_sys(const unsigned long long *p) {
	unsigned long long v;
	asm volatile("ld.relaxed.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
	return v;
}
__device__ __forceinline__ bool ll_ok(unsigned long long w, unsigned flag) { return (unsigned)(w >> 32) == flag; }
__device__ __forceinline__ float ll_val(unsigned long long w) { return __uint_as_float((unsigned)w); }
// Four consecutive elements (32 bytes); true when all four carry `flag`.
__device__ __forceinline__ bool ll_load4(const unsigned long long *p, unsigned flag, float4 *out) {
	const ulonglong2 a = ll_load2(p), b = ll_load2(p + 2);
	*out = make_float4(ll_val(a.x), ll_val(a.y), ll_val(b.x), ll_val(b.y));
	return ll_ok(a.x, flag) && ll_ok(a.y, flag) && ll_ok(b.x, flag) && ll_ok(b.y, flag);
}
// Back-off between polling rounds (keeps the pollers' L2 traffic away from the weight stream); a wait of ~4 s traps.
__device__ __forceinline__ void ll_backoff(int *round, long long *t0) {
	if (*round == 0) *t0 = clock64();
	else if (timed_out(*t0)) die(kFaultPoll, (unsigned)*round, 0);
	__nanosleep((unsigned)YALM_MEGA_LL_NS << min(*round, 4));
	++*round;
}
__device__ __forceinline__ float ll_wait1(const unsigned long long *p, unsigned flag) {
	unsigned long long w = ll_load1(p);
	if (!ll_ok(w, flag)) {
		int round = 0;
		long long t0 = 0;
		do {
			ll_backoff(&round, &t0);
			w = ll_load1(p);
		} while (!ll_ok(w, flag));
	}
	return ll_val(w);
}

__device__ __forceinline__ uint4 lds128(uint32_t addr) {
	uint4 v;
	asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(addr));
	return v;
}
__device__ __forceinline__ float4 lds128f(uint32_t addr) {
	float4 v;
	asm volatile("ld.shared.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(addr));
	return v;
}
// Packed fp32 FMA (sm_100 FFMA2): (a.x*b.x + c.x, a.y*b.y + c.y), each lane IEEE round-to-nearest like fmaf.
__device__ __forceinline__ float2 ffma2(float2 a, float2 b, float2 c) {
	unsigned long long d;
	asm("fma.rn.f32x2 %0, %1, %2, %3;"
	    : "=l"(d)
	    : "l"(*reinterpret_cast<unsigned long long *>(&a)), "l"(*reinterpret_cast<unsigned long long *>(&b)),
	      "l"(*reinterpret_cast<unsigned long long *>(&c)));
	return *reinterpret_cast<float2 *>(&d);
}
// acc += the EP weights of one 16-byte load times their x values, as EP/2 packed FMAs: acc.x collects the even
// elements, acc.y the odd ones (summed at the end of the row).  fp32 accumulate like the reference.
template <int DT>
__device__ __forceinline__ float2 fma16x2(uint4 w, const float4 *xv, float2 acc);
template <>
__device__ __forceinline__ float2 fma16x2<YALM_F32>(uint4 w, const float4 *xv, float2 acc) {
	acc = ffma2(make_float2(__uint_as_float(w.x), __uint_as_float(w.y)), make_float2(xv[0].x, xv[0].y), acc);
	acc = ffma2(make_float2(__uint_as_float(w.z), __uint_as_float(w.w)), make_float2(xv[0].z, xv[0].w), acc);
	return acc;
}
template <>
__device__ __forceinline__ float2 fma16x2<YALM_F16>(uint4 w, const float4 *xv, float2 acc) {
	using T = WTraits<YALM_F16>;
	acc = ffma2(T::h2(w.x), make_float2(xv[0].x, xv[0].y), acc);
	acc = ffma2(T::h2(w.y), make_float2(xv[0].z, xv[0].w), acc);
	acc = ffma2(T::h2(w.z), make_float2(xv[1].x, xv[1].y), acc);
	acc = ffma2(T::h2(w.w), make_float2(xv[1].z, xv[1].w), acc);
	return acc;
}
template <>
__device__ __forceinline__ float2 fma16x2<YALM_F8E4M3>(uint4 w, const float4 *xv, float2 acc) {
	using T = WTraits<YALM_F8E4M3>;
	const uint32_t ww[4] = {w.x, w.y, w.z, w.w};
#pragma unroll
	for (int i = 0; i < 4; ++i) {
		acc = ffma2(T::e2((uint16_t)(ww[i] & 0xffffu)), make_float2(xv[i].x, xv[i].y), acc);
		acc = ffma2(T::e2((uint16_t)(ww[i] >> 16)), make_float2(xv[i].z, xv[i].w), acc);
	}
	return acc;
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
			// K-split tiles are one bulk copy per row: below 4 KB per copy the single issuing thread, not
			// HBM, limits the ring (measured: 2 KB copies stream at 65 %), so prefer fewer rows per tile
			if (nk > 1 && KT * B < 4096 && R > (pair ? 2 : 1)) break;
			s->R = R, s->KT = KT, s->nk = nk;
			return true;
		}
	}
	return false;
}

// Optional phase profile (Args.prof != null): per CTA, per (layer, phase) kProfSlots clock64 values written by
// consumer thread 0: [0] phase entered (before the grid barrier) [1] barrier passed [2] input vector staged
// [3] tiles + epilogue done; [4] cycles warp 0 spent waiting for full ring stages (starvation = HBM-bound).
// Phase index of a launch -> (layer, kind of phase k, column half).  k: 0 qkv, 1 attention, 2 wo, 3 w1/w3, 4 w2, 5 LM head.
// phases per layer: qkv, attention, wo, w1/w3, then n_sub w2 sub-phases (dense 1; MoE one per selected expert)
__device__ __forceinline__ int ppl_of(const Args &a) { return 4 + a.n_sub; }
__device__ __forceinline__ void phase_of(int ph, const Args &a, int *l, int *k, int *sub) {
	const int ppl = ppl_of(a), j = ph % ppl;
	*l = min(ph / ppl, a.n_layers - 1);
	*sub = j >= 4 ? j - 4 : 0;
	*k = ph < a.n_layers * ppl ? min(j, 4) : 5;
}
constexpr int kProfSlots = 8, kProfPhases = 6; // phases: qkv, attn, wo, ffn, w2, lm(+argmax) (last layer slot only)

// Shared-memory layout: the small arrays sit at fixed offsets in front of the ring, so their addresses are
// compile-time constants instead of live registers:
//   rope table (64 float2) | red (16) | misc (32) | release counters (32) | full mbarriers | empty mbarriers | ring | vec
constexpr int kOffRope = 0, kOffRed = 512, kOffMisc = 576, kOffRel = 704, kOffFull = 832, kOffEmpty = kOffFull + 8 * kMaxStages;
// pipelined hb (below): full / empty mbarriers of the slice buffers | int [2][kHbMaxSlices]: hb elements this CTA still owes
// each slice, its share | int [kHbMaxSlices]: slice fetched (owned by the slice's duty warp)
constexpr int kOffHbFull = kOffEmpty + 8 * kMaxStages, kOffHbEmpty = kOffHbFull + 8 * kHbBufs;
constexpr int kOffHbCnt = (kOffHbEmpty + 8 * kHbBufs + 63) / 64 * 64, kOffHbIssued = kOffHbCnt + 8 * kHbMaxSlices;
// tensor-core GEMV (below): done / free mbarriers of the tiles' partial-sum slots
constexpr int kOffMmaDone = (kOffHbIssued + 4 * kHbMaxSlices + 63) / 64 * 64, kOffMmaFree = kOffMmaDone + 8 * kMmaSlots;
static_assert(kOffMmaFree + 8 * kMmaSlots <= kSmemHeader, "shared-memory header overflow");
constexpr int kOffStages = kSmemHeader; // 1024-byte aligned: swizzled tensor copies land in the stages
static_assert(kOffStages == kSmemHeader, "mega.h: kSmemHeader must match the kernel's shared-memory header");
__device__ __forceinline__ unsigned char *sm_base() {
	extern __shared__ __align__(1024) unsigned char smem_raw[];
	return smem_raw;
}
__device__ __forceinline__ float2 *sm_rope() { return reinterpret_cast<float2 *>(sm_base() + kOffRope); } // cos/sin of pos * freq_j
__device__ __forceinline__ float *sm_red() { return reinterpret_cast<float *>(sm_base() + kOffRed); }
__device__ __forceinline__ int *sm_misc() { return reinterpret_cast<int *>(sm_base() + kOffMisc); }
// sm_misc slots: [0] combine flag, [4, 4 + consumer warps) arg-max indices, [16..] parked context
constexpr int kMiscToken = 16, kMiscPos = 17, kMiscPark = 18;
constexpr int kMiscMmaTiles = 2; // tensor-core GEMV: the tile counter, parked across the attention unit loop
// MoE router record (written by consumer thread 0 in the w1/w3 staging step, polled by the producer warp):
// [kMiscRouter] = layer + 1 once [kMiscExpert + k] (ids) and [kMiscExpertW + k] (weights, as float bits) are set
constexpr int kMiscRouter = 23, kMiscExpert = 24, kMiscExpertW = 28;
static_assert(kMiscExpertW + kMegaMaxActive <= 32 && kMiscPark + 5 <= kMiscRouter, "sm_misc has 32 slots");
__device__ __forceinline__ int *sm_rel() { return reinterpret_cast<int *>(sm_base() + kOffRel); } // per-stage release counters
__device__ __forceinline__ uint32_t sm_full0() { return smem_u32(sm_base() + kOffFull); }
__device__ __forceinline__ uint32_t sm_empty0() { return smem_u32(sm_base() + kOffEmpty); }
__device__ __forceinline__ uint32_t sm_hbfull0() { return smem_u32(sm_base() + kOffHbFull); }
__device__ __forceinline__ uint32_t sm_hbempty0() { return smem_u32(sm_base() + kOffHbEmpty); }
__device__ __forceinline__ int *sm_hb_cnt() { return reinterpret_cast<int *>(sm_base() + kOffHbCnt); }
__device__ __forceinline__ volatile int *sm_hb_issued() { return reinterpret_cast<volatile int *>(sm_base() + kOffHbIssued); }
__device__ __forceinline__ uint32_t sm_mma_done0() { return smem_u32(sm_base() + kOffMmaDone); }
__device__ __forceinline__ uint32_t sm_mma_free0() { return smem_u32(sm_base() + kOffMmaFree); }
__device__ __forceinline__ unsigned char *sm_stages() { return sm_base() + kOffStages; }
__device__ __forceinline__ uint32_t sm_stages_u32() { return smem_u32(sm_base() + kOffStages); }

struct Ctx {
	long long *prof;  // this CTA's profile rows or null
#ifdef YALM_MEGA_TRACE
	long long *trace; // event trace (CTA 0 only) or null
#endif
#ifdef YALM_MEGA_TRACE
	long long wait_cycles;
#endif
	uint32_t vec_u32;
	float *vec;
	int cta, ncta, warp, lane, tid, S;
	uint32_t s_magic; // ceil(2^32 / S): idx / S == __umulhi(idx, s_magic) for every ring index of a launch
	uint32_t it; // ring index of the next tile of this role (same sequence for the producer and every consumer warp)
	int rot;     // CTA rotation of the current GEMV phase
	unsigned bar_gen;
	unsigned xgen; // cross-GPU barriers passed, earlier launches included (tensor parallelism)
	unsigned mt;   // tensor-core GEMV: tiles of this launch so far (same in every warp): slot and use count of a tile's partial sums
	int kv_sink, kv_pos, kv_len; // token and pos: sm_misc()[kMiscToken], [kMiscPos]
};

// Tile shape of a GEMV kind: planned on the host (tile_shape), read from the kernel-parameter constant bank
// (a dynamically indexed array inside Ctx would push the context into local memory).
__device__ __forceinline__ Shape shape_of(const Args &a, int kind) {
	Shape s;
	s.R = a.shape[kind][0], s.KT = a.shape[kind][1], s.nk = a.shape[kind][2], s.G = a.shape[kind][3];
	return s;
}

// Optional event trace of CTA 0 (Args.trace != null): trace[0] = event count, then (tag, clock64) pairs.
// tag = type << 56 | warp << 48 | ring index.  Types: 1 producer issued, 2 consumer wait begin, 3 wait end, 4 release,
// 5 phase stamp (index = layer * 64 + phase * 8 + slot).
// Compiled in only with -DYALM_MEGA_TRACE (the trace pointer and the starvation counter cost four registers that the
// attention unit loop does not have).
__device__ __forceinline__ void trace(const Ctx &c, int type, uint32_t idx) {
#ifdef YALM_MEGA_TRACE
	if (c.trace) {
		const unsigned long long k = atomicAdd(reinterpret_cast<unsigned long long *>(c.trace), 1ull);
		if (k < (unsigned long long)kTraceEvents) {
			c.trace[1 + 2 * k] = ((long long)type << 56) | ((long long)c.warp << 48) | (long long)idx;
			c.trace[2 + 2 * k] = clock64();
		}
	}
#endif
}

// (round, stage) of ring tile idx without an integer division
__device__ __forceinline__ void ring_pos(const Ctx &c, uint32_t idx, uint32_t *s, int *r) {
	const uint32_t q = __umulhi(idx, c.s_magic);
	*r = (int)q;
	*s = idx - q * (uint32_t)c.S;
}

__device__ __forceinline__ void stamp(Ctx &c, int layer, int phase, int k) {
	if (c.prof && c.tid == 0) {
		long long *row = c.prof + ((size_t)layer * kProfPhases + phase) * kProfSlots;
		row[k] = clock64();
		if (phase != 1 && (k == 2 || k == 3)) { // GEMV phases: slots 4 / 5 = the same two events on the GPU-wide clock (cross-CTA skew)
			unsigned long long g;
			asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(g));
			row[k == 3 ? 4 : 5] = (long long)g;
		}
#ifdef YALM_MEGA_TRACE
		if (k == 3 && phase != 1) row[4] = c.wait_cycles, c.wait_cycles = 0; // attention: slot 4 = combine probe
#endif
	}
	if (c.tid == 0) trace(c, 5, (uint32_t)(layer * 64 + phase * 8 + k));
}

// This CTA's share of a phase's G row groups: groups g_first, g_first + ncta, ... (NL of them).
__device__ __forceinline__ void my_groups(const Ctx &c, int G, int *g_first, int *NL) {
	*g_first = (c.cta - c.rot + c.ncta) % c.ncta;
	*NL = *g_first < G ? (G - *g_first + c.ncta - 1) / c.ncta : 0;
}

// Attention work split: KV head h belongs to the CTAs [ceil(h*ncta/n_kv), ceil((h+1)*ncta/n_kv)) — a CTA never
// spans two heads — and the head's nch row chunks are split contiguously over those CTAs.
struct AttnSplit {
	int kvh;    // this CTA's KV head
	int c0, nc; // first CTA and number of CTAs of the head
	int k0, nu; // this CTA's first chunk and number of chunks (units)
};
__host__ __device__ inline AttnSplit attn_split(int cta, int ncta, int n_kv, int nch) {
	AttnSplit s;
	s.kvh = (int)((unsigned)cta * n_kv / ncta);
	s.c0 = (s.kvh * ncta + n_kv - 1) / n_kv;
	s.nc = ((s.kvh + 1) * ncta + n_kv - 1) / n_kv - s.c0;
	const int r = cta - s.c0;
	s.k0 = (int)((unsigned)r * nch / s.nc);
	s.nu = (int)((unsigned)(r + 1) * nch / s.nc) - s.k0;
	return s;
}
// Does CTA c0 + r of a head with nc CTAs own a chunk?
__device__ __forceinline__ bool attn_has_unit(int r, int nc, int nch) {
	return nch >= nc || (unsigned)(r + 1) * nch / nc > (unsigned)r * nch / nc;
}

// ---- tensor-core GEMV: geometry (shared by the producer and the consumers) ---------------------------------------------
// A phase multiplies up to three matrices (QKV) by the staged vector.  CTA c owns a contiguous block of each matrix's rows
// (in units of 2 rows — RoPE and store pairs — or of 1 row of w1 AND of w3), cut into tiles of at most MT rows (MT = the
// tensor map's box rows, planned so that the largest block splits evenly: 28 rows -> 2 x 14); a tile's K is walked in
// stages of kMmaKCols columns.  Ring order: tile after tile, K stages inside; stage s of the phase belongs to warp s % W.
// Block of a matrix's rows owned by CTA `cta`: units of `unit` rows, split as evenly as integer division allows.
__host__ __device__ inline void mma_block(int cta, int ncta, int rows, int unit, int *row0, int *nrows) {
	const int n = rows / unit;
	const int u0 = (int)((long long)cta * n / ncta), u1 = (int)((long long)(cta + 1) * n / ncta);
	*row0 = u0 * unit, *nrows = (u1 - u0) * unit;
}
// Rows per tile (= box rows of the matrix's tensor map) for the largest block of `rows` rows over `grid` CTAs: at most `cap`,
// at least cap / 2, the value with the least over-fetch in a CTA's last tile (larger tiles on ties).
__host__ __device__ inline int mma_box_rows(int rows, int unit, int cap, int grid) {
	const int per = ((rows / unit + grid - 1) / grid) * unit;
	const int hi = cap < per ? cap : per, lo = unit > cap / 2 ? unit : cap / 2;
	int best = hi, waste = 1 << 30;
	for (int mt = hi; mt >= lo; mt -= unit) {
		const int w = (per + mt - 1) / mt * mt - per;
		if (w < waste) waste = w, best = mt;
	}
	return best;
}
struct MmaSub {
	int map;        // tensor map record
	int row0, nrows; // this CTA's rows of the matrix
	int MT, nt;     // rows per tile, tiles
	int goff;       // row offset of the matrix in the phase's output (q | k | v)
};
struct MmaPhase {
	MmaSub sub[3];
	int nsub, nkb, ntiles;
	bool ffn; // w1 rows -> tile rows 0-7, the same rows of w3 -> tile rows 8-15 (second box at +4096 bytes)
};
__device__ __forceinline__ MmaPhase mma_phase(const Args &a, int kind, int layer, int cta, int ncta) {
	MmaPhase p;
	p.nsub = 1, p.ffn = false, p.nkb = a.dim / kMmaKCols;
	const int mb = layer * kMmPerLayer;
	auto set = [&](int i, int map, int rows, int unit, int MT, int goff) {
		mma_block(cta, ncta, rows, unit, &p.sub[i].row0, &p.sub[i].nrows);
		p.sub[i].map = map;
		p.sub[i].MT = MT, p.sub[i].nt = (p.sub[i].nrows + MT - 1) / MT, p.sub[i].goff = goff;
	};
	p.sub[1].nt = p.sub[2].nt = 0;
	switch (kind) {
	case kQKV:
		p.nsub = 3;
		set(0, mb + kMmWq, a.q_dim, 2, a.mma_mt[0], 0);
		set(1, mb + kMmWk, a.kv_dim, 2, a.mma_mt[1], a.q_dim);
		set(2, mb + kMmWv, a.kv_dim, 2, a.mma_mt[1], a.q_dim + a.kv_dim);
		break;
	case kWO: set(0, mb + kMmWo, a.dim, 2, a.mma_mt[2], 0), p.nkb = a.q_dim / kMmaKCols; break;
	case kFFN: set(0, mb + kMmW1, a.hidden, 1, a.mma_mt[3], 0), p.ffn = true; break;
	case kW2: set(0, mb + kMmW2, a.dim, 2, a.mma_mt[4], 0), p.nkb = a.hidden / kMmaKCols; break;
	default: set(0, a.n_layers * kMmPerLayer, a.vocab, 2, a.mma_mt[5], 0); break;
	}
	p.ntiles = p.sub[0].nt + p.sub[1].nt + p.sub[2].nt;
	return p;
}

// ---- producer ---------------------------------------------------------------------------------------------
// Ring order inside a GEMV phase: this CTA's groups in blocks of kConsumerWarps; inside a block the K-split index
// is the slow axis and the group the fast one, so every consumer warp (one group each) has a tile in flight.
//
// One warp, all 32 lanes: a single thread needs ~1.5k cycles of dependent integer work per tile (measured: the
// ring starved), so the tile list is walked a batch at a time with lane l issuing tile (batch start + l).  A
// batch never exceeds the number of stages, so no lane waits on a stage another lane of the same batch fills.
//
// L2 look-ahead: with ring tile i of a phase the producer also issues an L2 prefetch for the tile `ahead`
// positions further down the CTA's tile sequence — crossing into the next phase — so HBM keeps streaming into L2
// while the consumers sit in a barrier / staging / attention-merge latency chain with the ring full, and the ring
// then refills from L2 at a multiple of the DRAM rate.  Invariant: on entry to a phase its first
// min(ahead, ntiles) tiles have been prefetched by the phase before it.
constexpr int kAttn = -1, kNone = -2;
struct PhaseDesc {
	int kind; // kQKV .. kLM, kAttn, or kNone (past the last phase)
	int ntiles;
	int nunits;          // look-ahead units: row groups (all K splits of R rows) or attention tiles
	uint32_t unit_bytes; // bytes of one unit
	uint32_t tile_bytes;
	// GEMV phases
	const char *m0, *m1, *m2;
	int rows0, rows1, K;
	Shape sh;
	int g_first, full_groups, full_tiles, nb_last;
	// attention
	const uint16_t *kc, *vc;
	int nch, u0;
	int rot_next; // CTA rotation after this phase
	// mixture of experts: the matrices hold n_experts stacked experts of `estride` bytes; which of them a tile reads is
	// known only once the router of `layer` has run (Producer::expert waits for it).  gpe: row groups per expert (w1/w3).
	int layer, sub, gpe;
	size_t estride;
	bool slice_major; // pipelined hb: w2 tiles in K-split-major order over ALL of the CTA's row groups
	bool mma;         // tensor-core GEMV: tiles = (tile, K stage) of mp, one 3-D tensor copy each
	MmaPhase mp;
	const char *tmaps;
};

struct Producer {
	Ctx &c;
	uint64_t policy;
	int batch; // min(stages, 32)
	int ahead;   // L2 look-ahead distance in tiles (0: off)
	bool steady; // false: only the next phase's prefix is fetched (during the phase-boundary bubble), nothing in-phase
	bool pf_tma; // L2 look-ahead through cp.async.bulk.prefetch.L2 (one request per range) instead of per-line prefetches

	template <int V>
	__device__ __forceinline__ PhaseDesc describe(const Args &a, int ph, int rot, int HD, int B) {
		PhaseDesc d;
		d.kind = kNone, d.ntiles = 0, d.nunits = 0, d.unit_bytes = kStageBytes, d.rot_next = rot, d.tile_bytes = 0;
		d.m0 = d.m1 = d.m2 = nullptr, d.rows0 = d.rows1 = d.K = 0, d.g_first = d.full_groups = d.full_tiles = d.nb_last = 0;
		d.kc = d.vc = nullptr, d.nch = 1, d.u0 = 0;
		d.sh.R = d.sh.KT = d.sh.nk = d.sh.G = 1;
		d.layer = 0, d.sub = 0, d.gpe = 0, d.estride = 0, d.slice_major = false, d.mma = false;
		const int n_phases = a.n_layers * ppl_of(a) + (a.mode != kHydrate ? 1 : 0);
		if (ph >= n_phases) return d;
		int l, k, sub;
		phase_of(ph, a, &l, &k, &sub);
		d.layer = l, d.sub = sub;
		const Layer &L = a.layers[l];
		if (k == 1) {
			const int C = kStageBytes / (4 * HD);
			d.kind = kAttn;
			d.nch = (c.kv_len + C - 1) / C;
			const AttnSplit sp = attn_split(c.cta, c.ncta, a.n_kv_heads, d.nch);
			d.u0 = sp.kvh * d.nch + sp.k0; // (head, chunk) as one index
			d.ntiles = sp.nu;
			d.nunits = d.ntiles;
			d.kc = L.kc, d.vc = L.vc;
			return d;
		}
		d.kind = kQKV, d.m0 = L.wq, d.m1 = L.wk, d.m2 = L.wv, d.rows0 = a.q_dim, d.rows1 = a.kv_dim, d.K = a.dim;
		switch (k) {
		case 2: d.kind = kWO, d.m0 = L.wo, d.rows0 = a.dim, d.K = a.q_dim; break;
		case 3: d.kind = kFFN, d.m0 = L.w1, d.m1 = L.w3, d.rows0 = a.hidden; break;
		case 4: d.kind = kW2, d.m0 = L.w2, d.rows0 = a.dim, d.K = a.hidden, d.slice_major = V == 1; break;
		case 5: d.kind = kLM, d.m0 = a.wcls, d.rows0 = a.vocab; break;
		default: break;
		}
		if (V == 2) { // tensor-core GEMV: the phase's stages follow from its geometry
			d.mma = true, d.mp = mma_phase(a, d.kind, l, c.cta, c.ncta), d.tmaps = static_cast<const char *>(a.tmaps);
			d.ntiles = d.nunits = d.mp.ntiles * d.mp.nkb;
			return d;
		}
		d.sh = shape_of(a, d.kind);
		if (a.n_experts > 0 && (k == 3 || k == 4)) {
			d.estride = (size_t)a.dim * a.hidden * B;
			d.gpe = k == 3 ? d.sh.G / a.n_sub : 0; // w1/w3: the phase's groups are n_sub experts x gpe groups
		}
		d.tile_bytes = (uint32_t)d.sh.R * d.sh.KT * B;
		d.g_first = (c.cta - rot + c.ncta) % c.ncta;
		const int NL = d.g_first < d.sh.G ? (d.sh.G - d.g_first + c.ncta - 1) / c.ncta : 0;
		d.ntiles = NL * d.sh.nk;
		d.nunits = NL, d.unit_bytes = d.tile_bytes * d.sh.nk;
		d.full_groups = NL / kConsumerWarps * kConsumerWarps, d.full_tiles = d.full_groups * d.sh.nk;
		d.nb_last = NL - d.full_groups;
		d.rot_next = (rot + d.sh.G) % c.ncta;
		if (k == 4 && sub + 1 < a.n_sub) d.rot_next = rot; // every w2 sub-phase keeps the row -> CTA map
		return d;
	}

	// Expert chosen for slot `slot` of layer `layer`: waits until the consumers' router step of that layer has published it
	// (the producer cannot run ahead of the router: which rows the w1/w3 and w2 phases read depends on it).
	__device__ __forceinline__ int expert(int layer, int slot) {
		volatile int *misc = sm_misc();
		if (misc[kMiscRouter] != layer + 1) {
			const long long t0 = clock64();
			while (misc[kMiscRouter] != layer + 1) {
				__nanosleep(32);
				if (timed_out(t0)) die(kFaultRouter, (unsigned)layer, (unsigned)slot);
			}
		}
		return misc[kMiscExpert + slot];
	}
	// Stage of ring tile idx is free (every earlier round consumed): expect `bytes` on its full barrier.
	__device__ __forceinline__ uint32_t acquire(uint32_t idx, uint32_t bytes, uint32_t *bar) {
		uint32_t s;
		int r;
		ring_pos(c, idx, &s, &r);
		if (r > 0) {
			// blocking wait on the stage's empty mbarrier (the hardware parks the thread: no issue slots burnt);
			// the release counter first rules out the one case a parity wait gets wrong (two rounds behind)
			if (ld_rel(sm_rel() + s) < r - 1) {
				const long long t0 = clock64();
				while (ld_rel(sm_rel() + s) < r - 1) {
					__nanosleep(64);
					if (timed_out(t0)) die(kFaultProducerRel, s, (unsigned)r);
				}
			}
			mbar_wait(sm_empty0() + 8 * s, (uint32_t)(r - 1) & 1u);
		}
		// the consumer's generic-proxy reads of the stage are ordered before the async-proxy refill
		asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
		*bar = sm_full0() + 8 * s;
		if (policy == 0) mbar_arrive(*bar); // debug & 8: no copies
		else mbar_expect_tx(*bar, bytes);
		trace(c, 1, idx);
		return sm_stages_u32() + s * kStageBytes;
	}
	// rows [row0, row0+nrows) x columns [kc*KT, (kc+1)*KT) of a (rows, K) matrix -> dst, tile rows packed
	// (bar == 0: pull the same bytes into L2 instead)
	__device__ __forceinline__ void rows_copy(uint32_t dst, uint32_t bar, const char *m, long row0, int nrows, int K, int KT,
	                                          int kc, int nk, int B) {
		const size_t row_bytes = (size_t)K * B;
		if (nk == 1 && KT == K) { // whole rows: one contiguous range
			if (bar) bulk_g2s(dst, m + row0 * row_bytes, (uint32_t)(nrows * row_bytes), bar, policy);
			else bulk_prefetch_l2(m + row0 * row_bytes, (uint32_t)(nrows * row_bytes), pf_tma);
		} else {
			const uint32_t tb = (uint32_t)KT * B;
			for (int r = 0; r < nrows; ++r) {
				const char *src = m + (row0 + r) * row_bytes + (size_t)kc * tb;
				if (bar) bulk_g2s(dst + r * tb, src, tb, bar, policy);
				else bulk_prefetch_l2(src, tb, pf_tma);
			}
		}
	}
	// Tile i of phase d -> ring index ring_idx (waits for its stage).  Returns the tile's look-ahead unit, or -1 if
	// the tile is not the first of its unit.
	template <int V>
	__device__ __forceinline__ int tile(const PhaseDesc &d, int i, uint32_t ring_idx, int HD, int max_seq, int B) {
		if (d.kind == kAttn) {
			const int u = d.u0 + i, kvh = u / d.nch, r0 = (u % d.nch) * (kStageBytes / (4 * HD));
			const int n = min(kStageBytes / (4 * HD), c.kv_len - r0);
			const uint32_t bytes = (uint32_t)n * HD * 2;
			const size_t off = ((size_t)kvh * max_seq + r0) * HD;
			uint32_t bar;
			const uint32_t dst = acquire(ring_idx, 2 * bytes, &bar);
			bulk_g2s(dst, d.kc + off, bytes, bar, policy);
			bulk_g2s(dst + kStageBytes / 2, d.vc + off, bytes, bar, policy);
			return i;
		}
		if (V == 2 && d.mma) {
			int T = i / d.mp.nkb;
			const int kb = i - T * d.mp.nkb;
			int m = 0;
			if (T >= d.mp.sub[0].nt) T -= d.mp.sub[0].nt, m = 1;
			if (m == 1 && T >= d.mp.sub[1].nt) T -= d.mp.sub[1].nt, m = 2;
			const MmaSub sb = m == 0 ? d.mp.sub[0] : m == 1 ? d.mp.sub[1] : d.mp.sub[2];
			const uint32_t box = 4u * (uint32_t)sb.MT * 128u;
			uint32_t bar;
			const uint32_t dst = acquire(ring_idx, d.mp.ffn ? 2 * box : box, &bar);
			const int row = sb.row0 + T * sb.MT;
			tma_load_3d(dst, d.tmaps + (size_t)sb.map * 128, 0, row, kb * (kMmaKCols / 64), bar, policy);
			if (d.mp.ffn) tma_load_3d(dst + 4096, d.tmaps + (size_t)(sb.map + (kMmW3 - kMmW1)) * 128, 0, row, kb * (kMmaKCols / 64), bar, policy);
			return -1;
		}
		int gl, kc;
		if (V == 1 && d.slice_major) {
			kc = i / d.nunits, gl = i - kc * d.nunits;
		} else if (i < d.full_tiles) {
			const int r = i % (kConsumerWarps * d.sh.nk);
			gl = i / (kConsumerWarps * d.sh.nk) * kConsumerWarps + r % kConsumerWarps, kc = r / kConsumerWarps;
		} else {
			const int r = i - d.full_tiles;
			gl = d.full_groups + r % d.nb_last, kc = r / d.nb_last;
		}
		int g = d.g_first + gl * c.ncta;
		size_t eoff = 0; // byte offset of the selected expert inside the stacked matrices
		if (d.estride) {
			const int slot = d.kind == kFFN ? g / d.gpe : d.sub;
			if (d.kind == kFFN) g -= slot * d.gpe;
			eoff = (size_t)expert(d.layer, slot) * d.estride;
		}
		uint32_t bar;
		const uint32_t dst = acquire(ring_idx, d.tile_bytes, &bar);
		if (d.kind == kFFN) {
			const int rh = d.sh.R / 2;
			rows_copy(dst, bar, d.m0 + eoff, (long)g * rh, rh, d.K, d.sh.KT, kc, d.sh.nk, B);
			rows_copy(dst + rh * d.sh.KT * B, bar, d.m1 + eoff, (long)g * rh, rh, d.K, d.sh.KT, kc, d.sh.nk, B);
		} else {
			long row = (long)g * d.sh.R;
			const char *m = d.m0 + eoff;
			if (d.kind == kQKV) {
				if (row >= d.rows0 + d.rows1) row -= d.rows0 + d.rows1, m = d.m2;
				else if (row >= d.rows0) row -= d.rows0, m = d.m1;
			}
			rows_copy(dst, bar, m, row, d.sh.R, d.K, d.sh.KT, kc, d.sh.nk, B);
		}
		return kc == 0 ? gl : -1;
	}
	// Pull look-ahead unit `unit` of phase d into L2: whole rows, one or two large requests (the copy engine is
	// bound by requests, not bytes, at the ring's 2 KB row pieces).
	__device__ __forceinline__ void prefetch_unit(const PhaseDesc &d, int unit, int HD, int max_seq, int B) {
		if (d.kind == kAttn) {
			const int u = d.u0 + unit, kvh = u / d.nch, r0 = (u % d.nch) * (kStageBytes / (4 * HD));
			const uint32_t bytes = (uint32_t)min(kStageBytes / (4 * HD), c.kv_len - r0) * HD * 2;
			const size_t off = ((size_t)kvh * max_seq + r0) * HD;
			bulk_prefetch_l2(d.kc + off, bytes, pf_tma);
			bulk_prefetch_l2(d.vc + off, bytes, pf_tma);
			return;
		}
		const int g = d.g_first + unit * c.ncta;
		const size_t row_bytes = (size_t)d.K * B;
		if (d.kind == kFFN) {
			const int rh = d.sh.R / 2;
			bulk_prefetch_l2(d.m0 + (size_t)g * rh * row_bytes, (uint32_t)(rh * row_bytes), pf_tma);
			bulk_prefetch_l2(d.m1 + (size_t)g * rh * row_bytes, (uint32_t)(rh * row_bytes), pf_tma);
		} else {
			long row = (long)g * d.sh.R;
			const char *m = d.m0;
			if (d.kind == kQKV) {
				if (row >= d.rows0 + d.rows1) row -= d.rows0 + d.rows1, m = d.m2;
				else if (row >= d.rows0) row -= d.rows0, m = d.m1;
			}
			bulk_prefetch_l2(m + row * row_bytes, (uint32_t)(d.sh.R * row_bytes), pf_tma);
		}
	}
	// Every phase of the token, in order (phase -1 is empty: it only fetches the prefix of phase 0).
	// `ahead` is in ring stages (kStageBytes); per phase it becomes a whole number of look-ahead units.
	template <int V>
	__device__ __forceinline__ void run(const Args &a, int HD, int B) {
		const int n_phases = a.n_layers * ppl_of(a) + (a.mode != kHydrate ? 1 : 0);
		const uint32_t ahead_bytes = (uint32_t)ahead * kStageBytes;
		PhaseDesc cur = describe<V>(a, n_phases, 0, HD, B); // kNone
		for (int ph = -1; ph < n_phases; ++ph) {
			const PhaseDesc nxt = describe<V>(a, ph + 1, cur.rot_next, HD, B);
			const int n = cur.ntiles, nu = cur.nunits;
			const int dist = (int)max(1u, ahead_bytes / cur.unit_bytes);                         // look-ahead of this phase, in its units
			const int npre = ahead > 0 ? min((int)max(1u, ahead_bytes / nxt.unit_bytes), nxt.nunits) : 0; // prefix of the next phase fetched here
			const int m = nu > 0 ? min(nu, dist) : 32; // the last m units of this phase share the next phase's prefix
			// a phase without tiles still walks one batch so that its lanes fetch the next prefix
			const int n_eff = n > 0 ? n : 32, nb = n > 0 ? batch : 32;
			for (int t0 = 0; t0 < n_eff; t0 += nb) {
				const int i = t0 + c.lane;
				if (c.lane < nb && i < n_eff) {
					const int unit = n > 0 ? tile<V>(cur, i, c.it + (uint32_t)i, HD, a.max_seq_len, B) : i;
					if (ahead > 0 && unit >= 0) {
						if (unit + dist < nu) {
							if (steady) prefetch_unit(cur, unit + dist, HD, a.max_seq_len, B);
						} else
							for (int j = unit - max(0, nu - dist); j < npre; j += m) prefetch_unit(nxt, j, HD, a.max_seq_len, B);
					}
				}
				__syncwarp();
			}
			c.it += (uint32_t)n;
			cur = nxt;
		}
	}
};

// ---- consumer helpers -------------------------------------------------------------------------------------
__device__ __forceinline__ float cons_sum(Ctx &c, float v) {
	v = warp_sum(v);
	cbar();
	if (c.lane == 0) sm_red()[c.warp] = v;
	cbar();
	float t = 0.f;
#pragma unroll
	for (int w = 0; w < kConsumerWarps; ++w) t += sm_red()[w];
	return t;
}

__device__ __forceinline__ void grid_barrier(Ctx &c, unsigned *sync) {
	++c.bar_gen;
	cbar(); // every consumer thread's writes of this phase happen-before thread 0's release below
	if (c.tid == 0) {
		const unsigned target = c.bar_gen * (unsigned)c.ncta;
		asm volatile("red.release.gpu.global.add.u32 [%0], 1;" ::"l"(sync) : "memory");
		unsigned v;
		asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(sync) : "memory");
		if (v < target) {
			const long long t0 = clock64();
			do {
				asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(sync) : "memory");
				if (timed_out(t0)) die(kFaultGridBarrier, v, target);
			} while (v < target);
		}
		asm volatile("fence.acq_rel.gpu;" ::: "memory");
	}
	cbar();
}

// Cross-GPU barrier of a tensor-parallel group: every CTA of every rank adds 1 to the arrival counter in EVERY rank's
// region (NVLink atomics, fire and forget) and waits until its own region has seen all world * ncta arrivals of this
// generation.  It is also the local grid barrier of the phase boundary.  The fence + relaxed atomics are a release at
// system scope: the partial rows this CTA pushed into the peers' slots are visible before its arrival is.
constexpr long long kTpTimeoutCycles = 120000000000ll; // ~60 s: ranks may enter a launch seconds apart; a lost peer traps
__device__ __forceinline__ void tp_barrier(Ctx &c, const Args &a) {
	++c.xgen;
	cbar();
	if (c.tid == 0) {
		asm volatile("fence.acq_rel.sys;" ::: "memory");
		for (int r = 0; r < a.tp_world; ++r)
			asm volatile("red.relaxed.sys.global.add.u32 [%0], 1;" ::"l"(a.tp_peer[r] + kTpOffBar) : "memory");
		const unsigned target = c.xgen * (unsigned)(c.ncta * a.tp_world);
		const char *mine = a.tp_peer[a.tp_rank] + kTpOffBar;
		unsigned v;
		asm volatile("ld.relaxed.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(mine) : "memory");
		if ((int)(v - target) < 0) {
			const long long t0 = clock64();
			do {
				asm volatile("ld.relaxed.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(mine) : "memory");
				if (timed_out(t0)) die(kFaultTpBarrier, v, target);
			} while ((int)(v - target) < 0);
		}
		asm volatile("fence.acq_rel.sys;" ::: "memory");
	}
	cbar();
}

// Wait until ring tile `idx` has landed; returns its shared-memory address.
__device__ __forceinline__ const unsigned char *wait_tile(Ctx &c, uint32_t idx, uint32_t *slot) {
	uint32_t s;
	int r;
	ring_pos(c, idx, &s, &r);
	const uint32_t par = (uint32_t)r & 1u;
	if (c.lane == 0) trace(c, 2, idx);
	if (ld_rel(sm_rel() + s) < r) { // an earlier round of this stage is still in use (rare)
		const long long t0 = clock64();
		while (ld_rel(sm_rel() + s) < r)
			if (timed_out(t0)) die(kFaultConsumerRel, s, (unsigned)r);
	}
#ifdef YALM_MEGA_TRACE
	if (c.prof && c.tid == 0) {
		const long long t0 = clock64();
		mbar_wait(sm_full0() + 8 * s, par);
		c.wait_cycles += clock64() - t0;
	} else
#endif
		mbar_wait(sm_full0() + 8 * s, par);
	if (c.lane == 0) trace(c, 3, idx);
	*slot = idx;
	return sm_stages() + (size_t)s * kStageBytes;
}
__device__ __forceinline__ uint32_t tile_u32(const Ctx &c, const unsigned char *tile) {
	return sm_stages_u32() + (uint32_t)(tile - sm_stages());
}
// Hand the stage of ring tile `idx` back to the producers.
__device__ __forceinline__ void release(Ctx &c, uint32_t idx) {
	__syncwarp();
	if (c.lane == 0) {
		uint32_t s;
		int r;
		ring_pos(c, idx, &s, &r);
		st_rel(sm_rel() + s, r + 1);
		mbar_arrive(sm_empty0() + 8 * s);
		trace(c, 4, idx); // release semantics: orders the warp's reads of the stage and the counter store
	}
}

// Four consecutive elements e .. e+3 (e % 4 == 0) of a phase's input vector -> shared memory.  SIMT path: fp32 in the
// permuted layout of xs_slot.  Tensor-core GEMV: an fp16 hi + lo pair (x = hi + lo to 2^-22) in the order of the mma B
// fragment: 16 bytes per (16-column k step, t) = [hi(2t, 2t+1) | hi(2t+8, 2t+9) | lo(2t, 2t+1) | lo(2t+8, 2t+9)], so that
// thread (g, t) of a warp fetches its fragment of column g & 1 (hi or lo) with one 8-byte load.
template <int EP, int V>
__device__ __forceinline__ void stage_store4(const Ctx &c, const Args &a, int e, float4 v) {
	if (V == 2) {
		const int ks = e >> 4, kk = e & 15; // kk in {0, 4, 8, 12}: elements kk, kk+1 -> t = (kk & 7) / 2, kk+2, kk+3 -> t + 1; same half
		char *base = reinterpret_cast<char *>(c.vec) + (size_t)(ks * 4 + ((kk & 7) >> 1)) * 16 + (kk >> 3) * 4;
		// hi is clamped to the fp16 range (lo then carries up to another 65504: |x| < 1.3e5 stays exact to 2^-22, beyond saturates)
		const __half2 h01 = __floats2half2_rn(fminf(fmaxf(v.x, -65504.f), 65504.f), fminf(fmaxf(v.y, -65504.f), 65504.f));
		const __half2 h23 = __floats2half2_rn(fminf(fmaxf(v.z, -65504.f), 65504.f), fminf(fmaxf(v.w, -65504.f), 65504.f));
		const float2 f01 = __half22float2(h01), f23 = __half22float2(h23);
		*reinterpret_cast<__half2 *>(base) = h01;
		*reinterpret_cast<__half2 *>(base + 8) = __floats2half2_rn(v.x - f01.x, v.y - f01.y);
		*reinterpret_cast<__half2 *>(base + 16) = h23;
		*reinterpret_cast<__half2 *>(base + 24) = __floats2half2_rn(v.z - f23.x, v.w - f23.y);
	} else {
		*reinterpret_cast<float4 *>(c.vec + xs_slot<EP>(e)) = v;
	}
}

// Input vector -> permuted shared layout (xs_slot).  embed_token < 0: fp32 vector in global memory (read through
// L2: other CTAs wrote it during this launch); >= 0: that row of the weight-typed embedding table (layer 0).
template <int DT>
__device__ __forceinline__ float4 load_in4(const Args &a, const float *src, int embed_token, int e) {
	if (embed_token < 0) return __ldcg(reinterpret_cast<const float4 *>(src + e));
	const size_t base = (size_t)embed_token * a.dim + e;
	const float sc = (DT == YALM_F8E4M3 && a.embed_scale) ? a.embed_scale[embed_token] : 1.0f;
	return make_float4(WTraits<DT>::to_float(a.embed, base) * sc, WTraits<DT>::to_float(a.embed, base + 1) * sc,
	                   WTraits<DT>::to_float(a.embed, base + 2) * sc, WTraits<DT>::to_float(a.embed, base + 3) * sc);
}

// The norm weights were last touched a token ago (14 GB of evict-first traffic since): they are loaded before the
// grid barrier, so that their DRAM latency hides behind the barrier instead of following it.
constexpr int kNormCache = 4;
__device__ __forceinline__ void load_norm_weights(const Ctx &c, const float *rms_w, int n, float4 (&wv)[kNormCache]) {
#pragma unroll
	for (int i = 0; i < kNormCache; ++i) {
		const int e = (c.tid + i * kCT) * 4;
		wv[i] = make_float4(0.f, 0.f, 0.f, 0.f);
		if (rms_w && e < n) wv[i] = *reinterpret_cast<const float4 *>(rms_w + e);
	}
}

// `wv`: this thread's norm weights (load_norm_weights), fetched before the grid barrier.
template <int DT, int V>
__device__ __forceinline__ void stage_vec(Ctx &c, const Args &a, const float *src, int n, const float *rms_w, int embed_token,
                                          const float4 (&wv)[kNormCache]) {
	constexpr int EP = WTraits<DT>::EP;
	constexpr int NC = kNormCache; // float4 per thread kept in registers between the two rmsnorm passes (n <= 4096)
	if (rms_w) {
		// reference formula and association order (src/infer.cpp:134-144): o = x * (1/sqrt(ss/n + eps)) * w
		float4 cache[NC];
		float ss = 0.f;
#pragma unroll
		for (int i = 0; i < NC; ++i) {
			const int e = (c.tid + i * kCT) * 4;
			cache[i] = make_float4(0.f, 0.f, 0.f, 0.f);
			if (e < n) cache[i] = load_in4<DT>(a, src, embed_token, e);
		}
#pragma unroll
		for (int i = 0; i < NC; ++i)
			ss += cache[i].x * cache[i].x + cache[i].y * cache[i].y + cache[i].z * cache[i].z + cache[i].w * cache[i].w;
		for (int e = (c.tid + NC * kCT) * 4; e < n; e += kCT * 4) {
			const float4 v = load_in4<DT>(a, src, embed_token, e);
			ss += v.x * v.x + v.y * v.y + v.z * v.z + v.w * v.w;
		}
		ss = cons_sum(c, ss);
		const float scale = 1.0f / sqrtf(ss / n + a.eps);
		// layer 0: the residual stream starts as the embedding row
		const bool seed_x = embed_token >= 0 && c.cta == 0;
		float *x_out = a.x;
#pragma unroll
		for (int i = 0; i < NC; ++i) {
			const int e = (c.tid + i * kCT) * 4;
			if (e < n) {
				const float4 v = cache[i], w = wv[i];
				stage_store4<EP, V>(c, a, e, make_float4(v.x * scale * w.x, v.y * scale * w.y, v.z * scale * w.z, v.w * scale * w.w));
				if (seed_x) *reinterpret_cast<float4 *>(x_out + e) = v;
			}
		}
		for (int e = (c.tid + NC * kCT) * 4; e < n; e += kCT * 4) {
			const float4 v = load_in4<DT>(a, src, embed_token, e);
			const float4 w = *reinterpret_cast<const float4 *>(rms_w + e);
			stage_store4<EP, V>(c, a, e, make_float4(v.x * scale * w.x, v.y * scale * w.y, v.z * scale * w.z, v.w * scale * w.w));
			if (seed_x) *reinterpret_cast<float4 *>(x_out + e) = v;
		}
	} else {
		int e = c.tid * 4;
		for (; e + 3 * kCT * 4 < n; e += 4 * kCT * 4) { // four independent L2 loads in flight per thread
			float4 v[4];
#pragma unroll
			for (int i = 0; i < 4; ++i) v[i] = __ldcg(reinterpret_cast<const float4 *>(src + e + i * kCT * 4));
#pragma unroll
			for (int i = 0; i < 4; ++i) stage_store4<EP, V>(c, a, e + i * kCT * 4, v[i]);
		}
		for (; e < n; e += kCT * 4)
			stage_store4<EP, V>(c, a, e, __ldcg(reinterpret_cast<const float4 *>(src + e)));
	}
	cbar();
}

// stage_vec for a flagged vector (see "flagged hand-over"): the loads ARE the phase barrier.  Every thread polls its
// own elements; round 0 issues all of a batch's loads back to back, later rounds only re-read what is still missing.
// The CTA barrier in front of the first shared-memory store also ends the previous phase for this CTA (its warps may
// still be reading c.vec in their tile loops).
template <int DT, int V>
__device__ __forceinline__ void stage_vec_ll(Ctx &c, const Args &a, const unsigned long long *src, int n, const float *rms_w,
                                             unsigned flag, const float4 (&wv)[kNormCache], long long *probe) {
	constexpr int EP = WTraits<DT>::EP;
	constexpr int NC = kNormCache; // wider batches (8, 12 float4 per thread) made ptxas spill all over the kernel
	float4 cache[NC];
	float ss = 0.f;
	// 148 CTAs x 352 threads polling the same few KB slow down the very stores they wait for (measured: ~5 us per
	// hand-over).  So one warp per CTA samples 32 words spread over the vector until they are there, everybody else
	// waits at the CTA barrier — which also ends the previous phase for this CTA: its warps may still be reading c.vec
	// in their tile loops — and only then all threads load (and verify: stragglers are re-read below).
	if (YALM_MEGA_LL_SAMPLE && c.warp == 0) {
		const int at = max(0, (int)((long)(c.lane + 1) * n / 32) - 1);
		unsigned long long w = ll_load1(src + at);
		if (!ll_ok(w, flag)) {
			const long long t0 = clock64();
			do {
				__nanosleep(40);
				if (timed_out(t0)) die(kFaultSample, 0, 0);
				w = ll_load1(src + at);
			} while (!ll_ok(w, flag));
		}
	}
	if (probe && c.tid == 0) probe[6] = clock64(); // profile: the sampled words are there
	cbar();
	if (YALM_MEGA_LL_FIRST && c.tid * 4 < n) {
		float4 t;
		if (!ll_load4(src + c.tid * 4, flag, &t)) {
			int round = 0;
			long long t0 = 0;
			do ll_backoff(&round, &t0);
			while (!ll_load4(src + c.tid * 4, flag, &t));
		}
	}
	for (int b0 = 0; b0 < n; b0 += NC * kCT * 4) { // one batch = NC float4 per thread (a single batch for n <= 5632); uniform trip count
		const int e0 = b0 + c.tid * 4;
		unsigned pending = 0;
#pragma unroll
		for (int i = 0; i < NC; ++i) {
			const int e = e0 + i * kCT * 4;
			cache[i] = make_float4(0.f, 0.f, 0.f, 0.f);
			if (e < n && !ll_load4(src + e, flag, &cache[i])) pending |= 1u << i;
		}
		if (pending) {
			int round = 0;
			long long t0 = 0;
			do {
				ll_backoff(&round, &t0);
#pragma unroll
				for (int i = 0; i < NC; ++i)
					if ((pending >> i & 1u) && ll_load4(src + e0 + i * kCT * 4, flag, &cache[i])) pending &= ~(1u << i);
			} while (pending);
		}
		if (rms_w) {
#pragma unroll
			for (int i = 0; i < NC; ++i)
				ss += cache[i].x * cache[i].x + cache[i].y * cache[i].y + cache[i].z * cache[i].z + cache[i].w * cache[i].w;
		} else {
#pragma unroll
			for (int i = 0; i < NC; ++i) {
				const int e = e0 + i * kCT * 4;
				if (e < n) stage_store4<EP, V>(c, a, e, cache[i]);
			}
		}
	}
	if (probe && c.tid == 0) probe[7] = clock64(); // profile: this thread's words are loaded and verified
	if (rms_w) {
		// reference formula and association order (src/infer.cpp:134-144): o = x * (1/sqrt(ss/n + eps)) * w
		ss = cons_sum(c, ss); // its CTA barriers come before any store to c.vec
		const float scale = 1.0f / sqrtf(ss / n + a.eps);
		const bool single = n <= NC * kCT * 4; // the only batch is still in registers, and so are its norm weights
		for (int b0 = 0; b0 < n; b0 += NC * kCT * 4) {
#pragma unroll
			for (int i = 0; i < NC; ++i) {
				const int e = b0 + (c.tid + i * kCT) * 4;
				if (e < n) {
					float4 v = cache[i], w = wv[i];
					if (!single) {
						ll_load4(src + e, flag, &v); // validated in the first pass
						w = *reinterpret_cast<const float4 *>(rms_w + e);
					}
					stage_store4<EP, V>(c, a, e, make_float4(v.x * scale * w.x, v.y * scale * w.y, v.z * scale * w.z, v.w * scale * w.w));
				}
			}
		}
	}
	cbar();
}

// ---- pipelined hb (w1/w3 -> w2, dense models; Args::w2_pipe) ---------------------------------------------------------
// hb is the largest vector that changes hands (hidden floats) and the phase that produces it is the longest of the layer:
// most of its rows are finished long before the phase is.  w2's K splits make natural slices — a w2 tile of K split kc needs
// hb[kc*KT, (kc+1)*KT) only.  So instead of staging all of hb behind the slowest CTA of w1/w3:
//   * the w1/w3 epilogue stores hb as plain fp32 in the permuted layout the w2 tile loop reads (hbp[xs_slot(row)]) and counts
//     the element down in a shared-memory counter of its slice (acq_rel at CTA scope); the thread that stores the CTA's LAST
//     element of a slice publishes the CTA's whole share with ONE GPU-scope fence + red on the slice's global completion
//     counter (a fence per element, even deferred to the thread's next epilogue, cost ~1.5 us each with the memory system
//     saturated: measured, the w1/w3 phase went from 31 to 45 us per layer);
//   * w2 walks its tiles K-split-major over ALL of the CTA's row groups (Producer::tile, slice_major), every warp carrying
//     the partial sums of its (at most kHbMaxGroups) groups across the slices, so a slice is used once and kHbBufs buffers of
//     one slice each replace the whole-hb buffer — shared memory that goes to the ring instead;
//   * slice s is pulled into buffer s % kHbBufs with one bulk copy by lane 0 of its duty warp (s % kConsumerWarps) as soon as
//     its counter is full and the buffer's previous slice has been released by all warps (acquire load, proxy fence,
//     cp.async.bulk completing on the buffer's mbarrier): the first kHbBufs slices while the CTA is still in the w1/w3 tile
//     loop, the others one slice ahead of their use;
//   * every warp waits for slice kc on the buffer's full mbarrier in front of its tiles of K split kc and arrives on the empty
//     mbarrier behind them (all warps, with or without tiles: every layer walks every slice, so use counts stay uniform).
// The w1/w3 -> w2 boundary has no staging step, no polling by 352 threads and no CTA barrier; only the LAST slice waits for
// the slowest CTA of w1/w3, and by then w2 has been streaming for microseconds.
__device__ __forceinline__ uint32_t hb_buf_u32(const Ctx &c, const Args &a, int b) {
	return sm_stages_u32() + (uint32_t)c.S * kStageBytes + (uint32_t)(a.hb_off + b * a.hb_slice) * 4;
}
// w1/w3 epilogue, after the thread's stores of one hb element of `slice`.
__device__ __forceinline__ void hb_count(const Args &a, int layer, int slice) {
	int *cnt = sm_hb_cnt() + slice;
	int old;
	if (a.debug & 32) asm volatile("atom.relaxed.cta.shared::cta.add.s32 %0, [%1], -1;" : "=r"(old) : "r"(smem_u32(cnt)) : "memory"); // timing experiment only
	else asm volatile("atom.acq_rel.cta.shared::cta.add.s32 %0, [%1], -1;" : "=r"(old) : "r"(smem_u32(cnt)) : "memory");
	if (old == 1) { // the CTA's share of the slice is complete, and every store to it happens-before this point
		asm volatile("fence.acq_rel.gpu;" ::: "memory");
		asm volatile("red.relaxed.gpu.global.add.u32 [%0], %1;" ::"l"(a.hb_cnt + (size_t)layer * kHbMaxSlices + slice), "r"(cnt[kHbMaxSlices]) : "memory");
	}
}
// hb elements of slice s that come from a CTA whose row groups are g_first, g_first + ncta, ... (NL of them, RH rows each);
// gps = row groups per slice.
__host__ __device__ inline int hb_share(int g_first, int NL, int ncta, int gps, int RH, int s) {
	const int lo = s * gps, hi = (s + 1) * gps; // groups below x owned by the CTA: 0 if x <= g_first, else ceil((x - g_first) / ncta) capped at NL
	int b_lo = lo <= g_first ? 0 : (lo - g_first + ncta - 1) / ncta, b_hi = hi <= g_first ? 0 : (hi - g_first + ncta - 1) / ncta;
	b_lo = b_lo < NL ? b_lo : NL, b_hi = b_hi < NL ? b_hi : NL;
	return (b_hi - b_lo) * RH;
}
// Start of the w1/w3 phase (in front of its staging step's CTA barriers): thread s < slices sets up slice s — how many hb
// elements this CTA's row groups g_first, g_first + ncta, ... (NL of them, RH rows each) contribute to it.
__device__ __forceinline__ void hb_setup(const Args &a, const Ctx &c) {
	const int s = c.tid;
	if (s >= a.shape[kW2][2]) return;
	const int RH = max(1, a.shape[kFFN][0] / 2);
	int g_first, NL;
	my_groups(c, a.shape[kFFN][3], &g_first, &NL);
	const int n = hb_share(g_first, NL, c.ncta, a.hb_slice / RH, RH, s);
	sm_hb_cnt()[s] = n, sm_hb_cnt()[kHbMaxSlices + s] = n;
	sm_hb_issued()[s] = 0;
}
// Lane 0 of slice s's duty warp: pull the slice into its buffer once it is complete and the buffer is free (`blocking`: wait
// for both).  `own`: called from the w1/w3 tile loop — do not even ask before this CTA's own share is in.
__device__ __forceinline__ void hb_fetch(const Ctx &c, const Args &a, int layer, int s, bool blocking, bool own) {
	if (own && *reinterpret_cast<volatile int *>(sm_hb_cnt() + s) != 0) return;
	const int idx = layer * a.shape[kW2][2] + s, b = idx % kHbBufs, r = idx / kHbBufs; // r-th use of buffer b in this launch
	if (r > 0) { // all warps have released the slice that used the buffer before
		const uint32_t eb = sm_hbempty0() + 8 * b, par = (uint32_t)(r - 1) & 1u;
		if (!mbar_try_wait(eb, par)) {
			if (!blocking) return;
			mbar_wait(eb, par);
		}
	}
	const unsigned *cn = a.hb_cnt + (size_t)layer * kHbMaxSlices + s;
	unsigned v = ld_acquire(cn);
	if (v != (unsigned)a.hb_slice) {
		if (!blocking) return;
		const long long t0 = clock64();
		do {
			__nanosleep(64);
			if (timed_out(t0)) die(kFaultHb, (unsigned)layer << 8 | (unsigned)s, v);
			v = ld_acquire(cn);
		} while (v != (unsigned)a.hb_slice);
	}
	// the bulk copy (async proxy) is ordered behind the acquire, and its shared-memory writes behind the warps' generic reads of
	// the buffer's previous slice (observed through the empty mbarrier)
	asm volatile("fence.proxy.async;" ::: "memory");
	const uint32_t bar = sm_hbfull0() + 8 * b, bytes = (uint32_t)a.hb_slice * 4;
	mbar_expect_tx(bar, bytes);
	bulk_g2s_plain(hb_buf_u32(c, a, b), a.hbp + (size_t)s * a.hb_slice, bytes, bar);
	sm_hb_issued()[s] = 1;
}

// Per-phase epilogue state.
struct Epi {
	int kind;
	const Layer *L;
	const float *scale0, *scale1, *scale2; // fp8 row scales per matrix
	int rows0, rows1;                      // QKV boundaries
	int sub, n_sub;                        // w2 sub-phase (MoE: one per selected expert; dense: 0 of 1)
	float ew;                              // MoE: gate weight of this sub-phase's expert (1 for dense models)
	int gpe;                               // MoE w1/w3: row groups per expert (0: dense)
	int layer;
	float best_v;                          // LM head: running arg-max of this lane's rows
	int best_i;
};

// acc2[r] += dot(row r of the tile at shared address `wa0` (R rows of KT columns), the input slice at `xa0`); both addresses
// already include the lane's 16-byte offset.
template <int DT, int R>
__device__ __forceinline__ void tile_fma(uint32_t wa, uint32_t xa, int KT, int lane, float2 (&acc2)[R]) {
	using T = WTraits<DT>;
	constexpr int EP = T::EP, CH = 32 * EP, Q = EP / 4;
	const int chunks = (KT + CH - 1) / CH, nfull = KT / CH;
	const int row_stride = KT * T::BYTES;
	int cc = 0;
	for (; cc + 2 <= nfull; cc += 2, wa += 1024, xa += 2 * CH * 4) { // all loads of two chunks, then the math
		uint4 w[2][R];
		float4 xv[2][Q];
#pragma unroll
		for (int u = 0; u < 2; ++u)
#pragma unroll
			for (int r = 0; r < R; ++r) w[u][r] = lds128(wa + r * row_stride + u * 512);
#pragma unroll
		for (int u = 0; u < 2; ++u)
#pragma unroll
			for (int q = 0; q < Q; ++q) xv[u][q] = lds128f(xa + (u * CH + q * 128) * 4);
#pragma unroll
		for (int u = 0; u < 2; ++u)
#pragma unroll
			for (int r = 0; r < R; ++r) acc2[r] = fma16x2<DT>(w[u][r], xv[u], acc2[r]);
	}
	for (; cc < chunks; ++cc, wa += 512, xa += CH * 4) { // odd full chunk and / or the ragged last one
		if (cc * CH + lane * EP < KT) {
			uint4 w[R];
			float4 xv[Q];
#pragma unroll
			for (int r = 0; r < R; ++r) w[r] = lds128(wa + r * row_stride);
#pragma unroll
			for (int q = 0; q < Q; ++q) xv[q] = lds128f(xa + q * 128 * 4);
#pragma unroll
			for (int r = 0; r < R; ++r) acc2[r] = fma16x2<DT>(w[r], xv, acc2[r]);
		}
	}
}

// Residual operand of the wo / w2 epilogue of row group g (lane = replica * R + row; tensor parallel: tp_collect adds it).
template <int R, bool TP>
__device__ __forceinline__ float load_resid(const Ctx &c, const Args &a, const Epi &ep, int g) {
	constexpr int NREP = YALM_MEGA_LL ? kLLRep : 1;
	const int kind = ep.kind, lane = c.lane;
	float resid = 0.f;
	if (!TP && (kind == kWO || kind == kW2) && lane < R * NREP) {
		// flagged hand-over: x after wo lives in ll_x[0] (complete: this CTA polled all of it when it staged w1/w3);
		// x before wo is the plain copy written behind w2 of the layer before, on the far side of a grid barrier
		if (kind == kW2 && ep.sub > 0) resid = __ldcg(a.w2_tmp + (long)g * R + lane % R); // x + the earlier experts' contributions
		else if (YALM_MEGA_LL && kind == kW2) resid = ll_val(ll_load1(a.ll_x + (size_t)(c.cta % kLLRep) * a.dim + (long)g * R + lane % R)); // the replica this CTA polled and verified
		else resid = __ldcg(a.x + (long)g * R + lane % R);
	}
	return resid;
}

// Epilogue of row group g: acc[r] = finished dot products of its R rows (valid in every lane); lane t handles output item t.
template <int DT, int R, bool TP, int V>
__device__ __forceinline__ void gemv_epilogue(Ctx &c, const Args &a, Epi &ep, int g, const float (&acc)[R], float resid, unsigned wflag) {
	constexpr bool LL = YALM_MEGA_LL;  // x: flagged words (tensor parallel: the finished rows, after the exchange at the end of the phase)
	constexpr bool LLH = YALM_MEGA_LL; // hb (and att_out): flagged too
	constexpr int NREP = LL ? kLLRep : 1;    // flagged vectors are written kLLRep times (lane -> replica): see kLLRep
	constexpr int NREPH = LLH ? kLLRep : 1;
	using T = WTraits<DT>;
	constexpr int EP = T::EP;
	const int kind = ep.kind, lane = c.lane;
	const int t = lane;
	if (kind == kQKV) {
		if (t < R / 2) {
			float v0 = acc[0], v1 = acc[R > 1 ? 1 : 0];
#pragma unroll
			for (int r = 2; r + 1 < R; r += 2)
				if (t == r / 2) v0 = acc[r], v1 = acc[r + 1];
			long row = (long)g * R + 2 * t;
			int which = 0;
			const float *sc = ep.scale0;
			if (row >= ep.rows0 + ep.rows1) which = 2, row -= ep.rows0 + ep.rows1, sc = ep.scale2;
			else if (row >= ep.rows0) which = 1, row -= ep.rows0, sc = ep.scale1;
			if (DT == YALM_F8E4M3 && sc) v0 *= sc[row], v1 *= sc[row + 1];
			v0 = clipf(v0, a.clip), v1 = clipf(v1, a.clip);
			const int hd = a.head_dim, j = (int)(row % hd);
			if (which < 2) { // interleaved-pair RoPE (src/infer.cpp:200-213), cos/sin from the per-token table
				const float2 cs = sm_rope()[j >> 1];
				const float r0 = v0 * cs.x - v1 * cs.y, r1 = v0 * cs.y + v1 * cs.x;
				v0 = r0, v1 = r1;
			}
			if (which == 0) {
				*reinterpret_cast<float2 *>(a.q + row) = make_float2(v0, v1);
			} else {
				__half *cache = reinterpret_cast<__half *>(which == 1 ? ep.L->kc : ep.L->vc);
				__half *dst = cache + ((size_t)(row / hd) * a.max_seq_len + c.kv_pos) * hd + j;
				*reinterpret_cast<__half2 *>(dst) = __floats2half2_rn(v0, v1);
			}
		}
	} else if (kind == kFFN) {
		constexpr int RH = R > 1 ? R / 2 : 1;
		if (t < RH * NREPH) { // every lane holds all sums: lane = replica * RH + row
			const int tr = t % RH;
			float gt = acc[0], up = acc[RH < R ? RH : 0];
#pragma unroll
			for (int r = 1; r < RH; ++r)
				if (tr == r) gt = acc[r], up = acc[RH + r < R ? RH + r : 0];
			const long row = (long)g * RH + tr; // MoE: row of hb = slot * hidden + i
			if (DT == YALM_F8E4M3 && ep.scale0) {
				long srow = row;
				if (ep.gpe) { // scales are stacked per expert like the weights
					const int slot = g / ep.gpe;
					srow = (long)sm_misc()[kMiscExpert + slot] * a.hidden + (row - (long)slot * a.hidden);
				}
				gt *= ep.scale0[srow], up *= ep.scale1[srow];
			}
			gt = a.act == YALM_ACT_GELU ? act_gelu(gt) : act_silu(gt);
			if (V == 1) { // pipelined hb: plain fp32 in the w2 tile loop's layout + the slice's completion count
				if (t < RH) {
					a.hb[row] = gt * up;
					a.hbp[xs_slot<EP>((int)row)] = gt * up;
					hb_count(a, ep.layer, (int)row / a.hb_slice);
				}
			} else {
				if (t < RH) a.hb[row] = gt * up;
				if (LLH) ll_store(a.ll_hb + (size_t)(t / RH) * a.hb_len + row, gt * up, wflag);
			}
		}
	} else if (TP) {
		// tensor parallel: the rows are partial sums (wo / w2: this rank's slice of the input dimension) or this
		// rank's slice of the vocabulary.  Every lane holds all R sums, so lane r hands the group's R values to
		// rank r over NVLink: flagged words into mailbox [set][this rank] (collected by the warp that owns the same
		// rows there: tp_collect below), or plain stores into the gathered logits.
		const long row0 = (long)g * R;
		float vv[R];
#pragma unroll
		for (int r = 0; r < R; ++r) {
			vv[r] = acc[r];
			if (DT == YALM_F8E4M3 && ep.scale0) vv[r] *= ep.scale0[row0 + r];
		}
		if (kind == kLM && t < R) {
			float v = vv[0];
#pragma unroll
			for (int r = 1; r < R; ++r)
				if (t == r) v = vv[r];
			const int row = a.vocab_base + (int)row0 + t;
			if (v > ep.best_v) ep.best_v = v, ep.best_i = row;
		}
		if (kind == kW2 && ep.n_sub > 1) { // MoE: gate weight, then on top of the earlier experts' sums (same lanes, earlier sub-phase)
#pragma unroll
			for (int r = 0; r < R; ++r) vv[r] = ep.sub > 0 ? fmaf(ep.ew, vv[r], __ldcg(a.w2_tmp + row0 + r)) : ep.ew * vv[r];
		}
		if (kind == kW2 && ep.sub + 1 < ep.n_sub) { // not the last expert: park the sums, nothing crosses the GPUs yet
			float v = vv[0];
#pragma unroll
			for (int r = 1; r < R; ++r)
				if (t == r) v = vv[r];
			if (t < R) a.w2_tmp[row0 + t] = v;
		} else if (lane < a.tp_world && kind == kLM) {
			float *dst = reinterpret_cast<float *>(a.tp_peer[lane] + kTpOffLogits) + a.vocab_base + row0;
			if (R == 4) *reinterpret_cast<float4 *>(dst) = make_float4(vv[0], vv[R > 1 ? 1 : 0], vv[R > 2 ? 2 : 0], vv[R > 3 ? 3 : 0]);
			else if (R == 2) *reinterpret_cast<float2 *>(dst) = make_float2(vv[0], vv[R > 1 ? 1 : 0]);
			else dst[0] = vv[0];
		} else if (lane < a.tp_world) {
			unsigned long long *dst = reinterpret_cast<unsigned long long *>(a.tp_peer[lane] + kTpOffMail) +
			                          (size_t)((kind == kW2 ? kTpMaxWorld : 0) + a.tp_rank) * kTpCapFloats + row0;
			if (R == 4) ll_store2_sys(dst, vv[0], vv[R > 1 ? 1 : 0], wflag), ll_store2_sys(dst + 2, vv[R > 2 ? 2 : 0], vv[R > 3 ? 3 : 0], wflag);
			else if (R == 2) ll_store2_sys(dst, vv[0], vv[R > 1 ? 1 : 0], wflag);
			else ll_store_sys(dst, vv[0], wflag);
		}
	} else if (t < (kind == kLM ? R : R * NREP)) { // wo / w2: lane = replica * R + row
		const int tr = t % R;
		float v = acc[0];
#pragma unroll
		for (int r = 1; r < R; ++r)
			if (tr == r) v = acc[r];
		const long row = (long)g * R + tr;
		if (DT == YALM_F8E4M3 && ep.scale0) v *= ep.scale0[row];
		if (kind == kLM) {
			a.logits[row] = v;
			if (v > ep.best_v) ep.best_v = v, ep.best_i = (int)row; // a lane's rows ascend: first maximum wins
		} else {
			// x += W v (src/infer.cpp:335-337); MoE: x += weight * (W2[e] hb_e), expert after expert (:377-383)
			const float xn = kind == kW2 && ep.n_sub > 1 ? fmaf(ep.ew, v, resid) : resid + v;
			if (kind == kW2 && ep.sub + 1 < ep.n_sub) {
				if (t < R) a.w2_tmp[row] = xn; // the next expert's sub-phase continues the row (same lane)
			} else {
				if (LL) ll_store(a.ll_x + (size_t)((kind == kW2 ? kLLRep : 0) + t / R) * a.dim + row, xn, wflag);
				if ((!LL || kind == kW2) && t < R) a.x[row] = xn;
			}
		}
	}
}

// Tensor parallel, behind wo and behind (the last sub-phase of) w2: finish this warp's rows.  Every rank's partial of group g
// has been (or is being) pushed into this rank's mailbox by the warp that owns g there; lane = rank * R + row polls its word,
// the sums run in rank order on top of the residual (the same bits on every rank), and the finished rows go out as local
// flagged words.
template <int R>
__device__ __forceinline__ void tp_collect(Ctx &c, const Args &a, int kind, int g_first, int NL, unsigned wflag, int ph) {
	constexpr int NREP = YALM_MEGA_LL ? kLLRep : 1;
	const int lane = c.lane, warp = c.warp;
	const unsigned long long *mail = reinterpret_cast<const unsigned long long *>(a.tp_peer[a.tp_rank] + kTpOffMail) +
	                                 (size_t)(kind == kW2 ? kTpMaxWorld : 0) * kTpCapFloats;
	const int W = a.tp_world;
	for (int b8 = 0; b8 < NL; b8 += kConsumerWarps) {
		if (warp >= min(kConsumerWarps, NL - b8)) break;
		const int g = g_first + (b8 + warp) * c.ncta;
		const int tr = lane % R;
		const long row = (long)g * R + tr;
		float part = 0.f;
		if (lane < W * R) {
			const unsigned long long *p = mail + (size_t)(lane / R) * kTpCapFloats + row;
			unsigned long long w = ll_load1_sys(p);
			if (!ll_ok(w, wflag)) {
				const long long t0 = clock64();
				do {
					__nanosleep(64);
					if (timed_out(t0)) die(kFaultTpMailbox, (unsigned)(lane / R) | (unsigned)ph << 8 | (unsigned)kind << 24, (unsigned)row);
					w = ll_load1_sys(p);
				} while (!ll_ok(w, wflag));
			}
			part = ll_val(w);
		}
		// residual: x before wo is the plain copy behind w2 of the layer before (far side of the attention grid barrier);
		// x before w2 is ll_x[0], complete in the replica this CTA verified when it staged w1/w3
		float xn = 0.f;
		if (lane < R * NREP)
			xn = kind == kW2 ? ll_val(ll_load1(a.ll_x + (size_t)(c.cta % kLLRep) * a.dim + row)) : __ldcg(a.x + row);
		for (int r = 0; r < W; ++r) xn += __shfl_sync(kFull, part, r * R + tr);
		if (lane < R * NREP) {
			ll_store(a.ll_x + (size_t)((kind == kW2 ? kLLRep : 0) + lane / R) * a.dim + row, xn, wflag);
			if (kind == kW2 && lane < R) a.x[row] = xn;
		}
	}
}

// All row groups of one GEMV phase owned by this CTA; the input vector is already staged in c.vec.
// Warp w owns groups w, w+8, ... of the CTA's list: it adds up all K-split tiles of the group, shuffle-reduces,
// and its first lanes run the epilogue.  Nothing here synchronises with the other warps.
template <int DT, int R, bool TP, int V>
__device__ __forceinline__ void gemv_consume(Ctx &c, const Args &a, Epi &ep, int ph) {
	static_assert(YALM_MEGA_LL || !TP, "tensor parallelism rides on the flagged hand-over");
	const unsigned wflag = ll_flag(a.seq, ph + 1); // flagged hand-over: the next phase polls for this
	using T = WTraits<DT>;
	constexpr int EP = T::EP, CH = 32 * EP;
	const Shape sh = shape_of(a, ep.kind);
	const int kind = ep.kind;
	const int chunks = (sh.KT + CH - 1) / CH;
	const int lane = c.lane, warp = c.warp;
	int g_first, NL;
	my_groups(c, sh.G, &g_first, &NL);
	const uint32_t base = c.it;
	const bool pipe_ffn = V == 1 && kind == kFFN; // pipelined hb: this phase produces the slices
	for (int b8 = 0; b8 < NL; b8 += kConsumerWarps) {
		const int nb = min(kConsumerWarps, NL - b8);
		if (warp >= nb) break;
		const int g = g_first + (b8 + warp) * c.ncta;
		if (pipe_ffn && warp < kHbBufs && warp < a.shape[kW2][2] && !(a.debug & 64)) { // duty warp of one of the first slices: fetch it once it is complete (never waits)
			if (lane == 0 && !sm_hb_issued()[warp]) hb_fetch(c, a, ep.layer, warp, false, true);
			__syncwarp();
		}
		// residual operand of the wo / w2 epilogue: issue the L2 read now, it is consumed after the tiles
		const float resid = load_resid<R, TP>(c, a, ep, g);
		float2 acc2[R];
#pragma unroll
		for (int r = 0; r < R; ++r) acc2[r] = make_float2(0.f, 0.f);
		for (int kc = 0; kc < sh.nk; ++kc) {
			uint32_t slot;
			const unsigned char *tile = wait_tile(c, base + (uint32_t)(b8 * sh.nk + kc * nb + warp), &slot);
			if (!(a.debug & 1))
				tile_fma<DT, R>(tile_u32(c, tile) + lane * 16, c.vec_u32 + (uint32_t)(kc * chunks * CH) * 4 + lane * 16, sh.KT, lane, acc2);
			release(c, slot);
		}
		float acc[R];
#pragma unroll
		for (int r = 0; r < R; ++r) acc[r] = warp_sum(acc2[r].x + acc2[r].y);
		gemv_epilogue<DT, R, TP, V>(c, a, ep, g, acc, resid, wflag);
	}
	if (TP && (kind == kWO || (kind == kW2 && ep.sub + 1 == ep.n_sub))) tp_collect<R>(c, a, kind, g_first, NL, wflag, ph);
	c.it = base + (uint32_t)(NL * sh.nk);
	if (!(kind == kW2 && ep.sub + 1 < ep.n_sub)) c.rot = (c.rot + sh.G) % c.ncta; // w2 sub-phases share one row -> CTA map
}

// w2 with pipelined hb (see "pipelined hb"): K-split-major over all of this CTA's row groups.  Warp w owns groups w, w + W, ...
// (J <= kHbMaxGroups of them) and keeps their partial sums in registers across the slices; ring tile of (K split kc, group gl)
// = base + kc * NL + gl.  The group loop is a run-time loop over ONE copy of the tile code: the register arrays rotate by one
// group per iteration (plain moves), so the current group is always index 0.
template <int DT, int R, bool TP>
__device__ __forceinline__ void w2_pipe_consume(Ctx &c, const Args &a, Epi &ep, int ph) {
	constexpr int W = kConsumerWarps, JM = kHbMaxGroups;
	const unsigned wflag = ll_flag(a.seq, ph + 1);
	const Shape sh = shape_of(a, kW2);
	const int lane = c.lane, warp = c.warp;
	int g_first, NL;
	my_groups(c, sh.G, &g_first, &NL);
	const uint32_t base = c.it;
	const int J = warp < NL ? (NL - warp + W - 1) / W : 0;
	float2 acc2[JM][R];
	float resid[JM];
#pragma unroll
	for (int j = 0; j < JM; ++j) {
		resid[j] = j < J ? load_resid<R, TP>(c, a, ep, g_first + (warp + W * j) * c.ncta) : 0.f;
#pragma unroll
		for (int r = 0; r < R; ++r) acc2[j][r] = make_float2(0.f, 0.f);
	}
	auto rotate = [&]() {
#pragma unroll
		for (int r = 0; r < R; ++r) {
			const float2 t = acc2[0][r];
#pragma unroll
			for (int j = 0; j + 1 < JM; ++j) acc2[j][r] = acc2[j + 1][r];
			acc2[JM - 1][r] = t;
		}
		const float t = resid[0];
#pragma unroll
		for (int j = 0; j + 1 < JM; ++j) resid[j] = resid[j + 1];
		resid[JM - 1] = t;
	};
	// lane 0: fetch slice s if this warp is its duty warp and it has not been fetched yet
	auto duty = [&](int s, bool blocking) {
		if (s < sh.nk && s % W == warp) {
			if (lane == 0 && !sm_hb_issued()[s]) hb_fetch(c, a, ep.layer, s, blocking, false);
			__syncwarp();
		}
	};
	for (int kc = 0; kc < sh.nk; ++kc) {
		const int idx = ep.layer * sh.nk + kc, b = idx % kHbBufs;
		duty(kc, true);
#pragma unroll 1
		for (int s = kc + 1; s < kc + kHbBufs; ++s) duty(s, false); // ahead: slice s's buffer is free once every warp has left slice s - kHbBufs
		const bool probe = c.prof && c.tid == 0 && kc == sh.nk - 1; // profile: how long the LAST slice keeps warp 0 waiting
		if (probe) c.prof[((size_t)ep.layer * kProfPhases + 4) * kProfSlots + 7] = clock64();
		mbar_wait(sm_hbfull0() + 8 * b, (uint32_t)(idx / kHbBufs) & 1u);
		if (probe) c.prof[((size_t)ep.layer * kProfPhases + 4) * kProfSlots + 6] = clock64();
		const uint32_t xa = hb_buf_u32(c, a, b) + lane * 16;
#pragma unroll 1
		for (int j = 0; j < JM; ++j) {
			if (j < J) {
				uint32_t slot;
				const unsigned char *tile = wait_tile(c, base + (uint32_t)(kc * NL + warp + W * j), &slot);
				if (!(a.debug & 1)) tile_fma<DT, R>(tile_u32(c, tile) + lane * 16, xa, sh.KT, lane, acc2[0]);
				release(c, slot);
			}
			rotate();
		}
#pragma unroll 1
		for (int s = kc + 1; s < kc + kHbBufs; ++s) duty(s, false); // second chance: the slower warps have moved on meanwhile
		__syncwarp();
		if (lane == 0) mbar_arrive(sm_hbempty0() + 8 * b);
	}
#pragma unroll 1
	for (int j = 0; j < JM; ++j) {
		if (j < J) {
			float acc[R];
#pragma unroll
			for (int r = 0; r < R; ++r) acc[r] = warp_sum(acc2[0][r].x + acc2[0][r].y);
			gemv_epilogue<DT, R, TP, 1>(c, a, ep, g_first + (warp + W * j) * c.ncta, acc, resid[0], wflag);
		}
		rotate();
	}
	if (TP) tp_collect<R>(c, a, kW2, g_first, NL, wflag, ph);
	c.it = base + (uint32_t)(NL * sh.nk);
	c.rot = (c.rot + sh.G) % c.ncta;
}

// ---- tensor-core GEMV (fp16 dense models; Args::mma) -------------------------------------------------------------------
// The SIMT tile loop above is at its floor — 63 instructions per 2 KB of fp16 weights, half of them fp16 -> fp32
// conversions — and with three resident warps per scheduler it runs at 1.5x the rate HBM delivers, so every stall of a
// phase boundary is paid in full.  Here the multiply-accumulate goes to the tensor cores:
//   * a ring stage is a 16-row x 256-column tile of the row-major weight matrix, fetched by ONE 3-D tensor copy (64 columns
//     x MT rows x 4 column blocks, SWIZZLE_128B): it lands as [column block][row][128 bytes] with the 16-byte chunks of every
//     row XOR-ed with address bits 7-9, which makes the eight row addresses of an ldmatrix conflict-free;
//   * x sits in shared memory as an fp16 hi + lo pair in B-fragment order (stage_store4): B column 0 = hi, column 1 = lo, so
//     D[row][0] + D[row][1] = W x to 2^-22 relative, fp32 accumulation in the tensor core; the other six columns are ignored;
//   * per 16-column k step: one ldmatrix.x4 (A), one 8-byte load (B), one mma.sync.m16n8k16 — 48 + addressing instructions per
//     8 KB stage instead of 252;
//   * the K stages of a tile go round the consumer warps (stage s -> warp s % W), every warp keeps its partial D in four
//     registers while it stays on the tile, then parks the 16 row sums in slot (tile % kMmaSlots) and arrives on the slot's
//     `done` mbarrier; the tile's owner warp (tile % W) adds the W partials in warp order — deterministic — and runs the same
//     pair-wise epilogue as the SIMT path (gemv_epilogue<2>), then frees the slot.
// Measured (Mistral-7B fp16, 4k context, one B200): tokens identical to the SIMT path and the reference; without
// any copies a token takes 2.21 ms instead of 2.29 (452 vs 436 tok/s) — but the stream of 3-D tensor copies (7 KB in an 8 KB
// stage, 512 contiguous bytes per row) runs 7 % below the 1-D row pieces (340 vs 367 tok/s with the arithmetic skipped on
// both), so the whole token is 311 vs 326 tok/s: the path is opt-in (YALM_MEGA_MMA=1), not the default.
// Rows past the CTA's block inside its last tile belong to the next CTA, which computes them itself: they are fetched twice
// (+1 % of the bytes at MT = 14) and masked here.  w1/w3: tile rows 0-7 are rows of w1, rows 8-15 the same rows of w3, so the
// gate and the up value of an hb element meet in one thread.
// Epilogue of one tile, lane-parallel (fp16 dense models on one GPU: no row scales, no experts, no mailboxes): y = the row
// sums in lanes 0-15 (w1/w3: gate rows in lanes 0-7, up rows in lanes 8-15), vr valid rows from row `row_t` of matrix m.
// Same arithmetic as gemv_epilogue, one item per lane instead of one call per row pair.  `resid`: see the call site.
__device__ __forceinline__ void mma_epilogue(Ctx &c, const Args &a, Epi &ep, int m, int row_t, int vr, float y, float resid, unsigned wflag) {
	const int lane = c.lane, kind = ep.kind;
	if (kind == kFFN) { // lane = replica * 8 + element
		const int p = lane & 7, rp = lane >> 3;
		const float gate = __shfl_sync(kFull, y, p), up = __shfl_sync(kFull, y, 8 + p);
		if (p < vr && rp < kLLRep) {
			const float gt = a.act == YALM_ACT_GELU ? act_gelu(gate) : act_silu(gate);
			const long row = row_t + p;
			if (rp == 0) a.hb[row] = gt * up;
			ll_store(a.ll_hb + (size_t)rp * a.hb_len + row, gt * up, wflag);
		}
	} else if (kind == kQKV) { // lane = row pair
		float v0 = __shfl_sync(kFull, y, (2 * lane) & 31), v1 = __shfl_sync(kFull, y, (2 * lane + 1) & 31);
		if (2 * lane < vr) {
			const int row = row_t + 2 * lane, hd = a.head_dim, j = row % hd; // row inside wq (m 0) / wk (1) / wv (2)
			v0 = clipf(v0, a.clip), v1 = clipf(v1, a.clip);
			if (m < 2) { // interleaved-pair RoPE (src/infer.cpp:200-213), cos/sin from the per-token table
				const float2 cs = sm_rope()[j >> 1];
				const float r0 = v0 * cs.x - v1 * cs.y, r1 = v0 * cs.y + v1 * cs.x;
				v0 = r0, v1 = r1;
			}
			if (m == 0) {
				*reinterpret_cast<float2 *>(a.q + row) = make_float2(v0, v1);
			} else {
				__half *cache = reinterpret_cast<__half *>(m == 1 ? ep.L->kc : ep.L->vc);
				*reinterpret_cast<__half2 *>(cache + ((size_t)(row / hd) * a.max_seq_len + c.kv_pos) * hd + j) = __floats2half2_rn(v0, v1);
			}
		}
	} else if (kind == kLM) { // lane = row
		if (lane < vr) {
			const int row = row_t + lane;
			a.logits[row] = y;
			if (y > ep.best_v) ep.best_v = y, ep.best_i = row; // a lane's rows ascend: first maximum wins
		}
	} else { // wo / w2: x += W v (src/infer.cpp:335-337); lane = (replica & 1) * 16 + row, two passes over the replicas
		const int r = lane & 15;
		const float xn = resid + __shfl_sync(kFull, y, r);
		if (r < vr) {
			const long row = row_t + r;
			unsigned long long *dst = a.ll_x + (size_t)(kind == kW2 ? kLLRep : 0) * a.dim + row;
#pragma unroll
			for (int rp = lane >> 4; rp < kLLRep; rp += 2) ll_store(dst + (size_t)rp * a.dim, xn, wflag);
			if (kind == kW2 && lane < 16) a.x[row] = xn;
		}
	}
}

template <int DT, bool TP>
__device__ __forceinline__ void mma_consume(Ctx &c, const Args &a, Epi &ep, int ph) {
	constexpr int W = kConsumerWarps;
	const unsigned wflag = ll_flag(a.seq, ph + 1);
	const int lane = c.lane, warp = c.warp, g8 = lane >> 2, t4 = lane & 3;
	const MmaPhase mp = mma_phase(a, ep.kind, ep.layer, c.cta, c.ncta);
	const uint32_t base = c.it;
	float *part = c.vec + (a.vec_floats - kMmaSlots * W * 16); // [slot][warp][16 rows]
	// ldmatrix geometry of this lane: matrix lane / 8 = a0..a3 = (rows 0-7 | 8-15) x (k 0-7 | 8-15); it supplies one row address
	const int mi = lane >> 3, arow = (lane & 7) + ((mi & 1) << 3), kch = mi >> 1;
	const uint32_t xb = c.vec_u32 + (uint32_t)t4 * 16 + (uint32_t)(g8 & 1) * 8; // B fragment of column g8 & 1 (hi | lo)
	int T = 0;
	for (int m = 0; m < mp.nsub; ++m) {
		const MmaSub sb = m == 0 ? mp.sub[0] : m == 1 ? mp.sub[1] : mp.sub[2];
		const uint32_t rstride = (uint32_t)sb.MT * 128; // one column block of the box
		const uint32_t roff = mp.ffn ? (arow < 8 ? 0u : 4096u) + (uint32_t)(arow & 7) * 128 : (uint32_t)arow * 128;
		for (int t = 0; t < sb.nt; ++t, ++T) {
			float acc[4][4]; // four independent accumulator sets (k step % 4): a single set is one dependent chain of 16 mma per stage
#pragma unroll
			for (int q = 0; q < 4; ++q)
#pragma unroll
				for (int i = 0; i < 4; ++i) acc[q][i] = 0.f;
			const int s0 = T * mp.nkb;
			for (int kb = (warp - s0 % W + W) % W; kb < mp.nkb; kb += W) {
				uint32_t slot;
				const unsigned char *tile = wait_tile(c, base + (uint32_t)(s0 + kb), &slot);
				if (!(a.debug & 1)) {
					const uint32_t tb = tile_u32(c, tile) + roff, xk = xb + (uint32_t)kb * (kMmaKCols / 16 * 64);
#pragma unroll
					for (int j = 0; j < 4; ++j) {
						const uint32_t ra = tb + j * rstride, sw = (ra >> 7) & 7u;
#pragma unroll
						for (int q = 0; q < 4; ++q) { // k step 4 j + q: 16-byte chunks 2 q (k 0-7) and 2 q + 1 (k 8-15) of column block j
							uint32_t A[4];
							ldmatrix_x4(A, ra + ((((uint32_t)(2 * q + kch)) ^ sw) << 4));
							const uint2 B = lds64(xk + (uint32_t)(4 * j + q) * 64);
							mma16816(acc[q], A, B.x, B.y);
						}
					}
				}
				release(c, slot);
			}
			// park this warp's 16 row sums (zero if it had no stage of the tile) and report
			const unsigned Tg = c.mt + (unsigned)T, ps = Tg % kMmaSlots, use = Tg / kMmaSlots;
			if (use > 0) mbar_wait(sm_mma_free0() + 8 * ps, (use - 1) & 1u); // the slot's previous tile has been summed
			float *pp = part + (ps * W + warp) * 16;
			if (t4 == 0) { // columns 0 (W hi) + 1 (W lo) of rows g8 and g8 + 8
				pp[g8] = ((acc[0][0] + acc[0][1]) + (acc[1][0] + acc[1][1])) + ((acc[2][0] + acc[2][1]) + (acc[3][0] + acc[3][1]));
				pp[g8 + 8] = ((acc[0][2] + acc[0][3]) + (acc[1][2] + acc[1][3])) + ((acc[2][2] + acc[2][3]) + (acc[3][2] + acc[3][3]));
			}
			__syncwarp();
			if (lane == 0) mbar_arrive(sm_mma_done0() + 8 * ps);
			if (Tg % W != (unsigned)warp) continue;
			// ---- owner of the tile: sum the partials in warp order, then the lane-parallel epilogue ----------------------
			const int row_t = sb.row0 + t * sb.MT, vr = min(sb.MT, sb.nrows - t * sb.MT);
			// residual operand of wo / w2 (lane & 15 = row), in flight across the wait: x after wo lives in ll_x[0] (the replica
			// this CTA verified when it staged w1/w3), x before wo is the plain copy written behind w2 of the layer before
			float resid = 0.f;
			if ((ep.kind == kWO || ep.kind == kW2) && (lane & 15) < vr) {
				const long row = row_t + (lane & 15);
				resid = ep.kind == kW2 ? ll_val(ll_load1(a.ll_x + (size_t)(c.cta % kLLRep) * a.dim + row)) : __ldcg(a.x + row);
			}
			mbar_wait(sm_mma_done0() + 8 * ps, use & 1u);
			float y = 0.f;
			if (lane < 16) {
				const float *q = part + ps * W * 16 + lane;
#pragma unroll
				for (int w = 0; w < W; ++w) y += q[w * 16];
			}
			__syncwarp();
			if (lane == 0) mbar_arrive(sm_mma_free0() + 8 * ps);
			mma_epilogue(c, a, ep, m, row_t, vr, y, resid, wflag);
		}
	}
	c.mt += (unsigned)mp.ntiles;
	c.it = base + (uint32_t)(mp.ntiles * mp.nkb);
}

// Kernel variant V (compile time: the default kernel carries none of the optional paths — with all of them in one kernel it
// had 60 % more code and its instruction-cache hit rate fell from 94 to 91 %): 0 = flagged hand-over everywhere, SIMT tile
// loops; 1 = pipelined hb; 2 = tensor-core GEMV.  Same-box A/B of variant 0 against the kernel before these paths existed:
// 329.3 / 329.8 vs 329.4 / 328.7 tok/s.
template <int DT, bool TP, int V>
__device__ __forceinline__ void gemv_phase(Ctx &c, const Args &a, Epi &ep, int ph) {
	if constexpr (V == 2) {
		mma_consume<DT, TP>(c, a, ep, ph);
	} else {
		const int R = a.shape[ep.kind][0];
		if (V == 1 && ep.kind == kW2) { // plan(): R <= 2
			if (R == 2) w2_pipe_consume<DT, 2, TP>(c, a, ep, ph);
			else w2_pipe_consume<DT, 1, TP>(c, a, ep, ph);
		} else if (R == 4) gemv_consume<DT, 4, TP, V>(c, a, ep, ph);
		else if (R == 2) gemv_consume<DT, 2, TP, V>(c, a, ep, ph);
		else gemv_consume<DT, 1, TP, V>(c, a, ep, ph);
	}
}

// ---- attention --------------------------------------------------------------------------------------------
__device__ __forceinline__ void unpack8(uint4 v, float (&f)[8]) {
	float2 a = __half22float2(*reinterpret_cast<const __half2 *>(&v.x));
	float2 b = __half22float2(*reinterpret_cast<const __half2 *>(&v.y));
	float2 c = __half22float2(*reinterpret_cast<const __half2 *>(&v.z));
	float2 d = __half22float2(*reinterpret_cast<const __half2 *>(&v.w));
	f[0] = a.x, f[1] = a.y, f[2] = b.x, f[3] = b.y, f[4] = c.x, f[5] = c.y, f[6] = d.x, f[7] = d.y;
}

// Last CTA of a KV head: combine the participating CTAs' partials (acc, max, sum) in CTA order into att_out.
// Runs on 1 CTA per head and layer, so it is kept out of line: its registers do not weigh on the main loops.
template <int HD, int G>
__device__ __forceinline__ void combine_head(const float *ws_acc, const float *ws_ml, float *att_out, float *wts, float *scratch,
                                          int kvh, int ncta, int nch, int b_lo, int b_hi, long long *probe,
                                          unsigned long long *ll_att = nullptr, unsigned wflag = 0, int n_q = 0) {
	constexpr int W = kConsumerWarps;
	constexpr int kInFlight = 20;  // independent L2 reads per thread and output
	const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
	const int nc = b_hi - b_lo + 1;
	for (int g = warp; g < G; g += W) {
		const float2 *mlp = reinterpret_cast<const float2 *>(ws_ml) + (size_t)kvh * ncta * G + g;
		// the first 32 CTAs of the range stay in registers: one L2 round trip in the common case
		const int b0 = b_lo + lane;
		const bool p0 = b0 <= b_hi && attn_has_unit(lane, nc, nch);
		float2 ml0 = make_float2(-INFINITY, 0.f);
		if (p0) ml0 = __ldcg(mlp + (size_t)b0 * G);
		float M = ml0.x;
		for (int b = b0 + 32; b <= b_hi; b += 32)
			if (attn_has_unit(b - b_lo, nc, nch)) M = fmaxf(M, __ldcg(mlp + (size_t)b * G).x);
		M = warp_max(M);
		float ls = 0.f;
		if (b0 <= b_hi) {
			const float w = !p0 ? -1.f : ml0.x == -INFINITY ? 0.f : expf(ml0.x - M);
			if (p0) ls = w * ml0.y;
			wts[b0 * G + g] = w;
		}
		for (int b = b0 + 32; b <= b_hi; b += 32) {
			float w = -1.f;
			if (attn_has_unit(b - b_lo, nc, nch)) {
				const float2 ml = __ldcg(mlp + (size_t)b * G);
				w = ml.x == -INFINITY ? 0.f : expf(ml.x - M);
				ls += w * ml.y;
			}
			wts[b * G + g] = w;
		}
		ls = warp_sum(ls);
		if (lane == 0) wts[ncta * G + g] = ls;
	}
	cbar();
	if (probe && tid == 0) probe[0] = clock64();
	// two outputs per thread, kInFlight independent L2 reads each, summed in CTA order.  Branch-free: CTAs of the
	// range that hold no unit of this head (weight -1) are read at b_hi's address (the last CTA of a head always
	// owns a unit) and weighted 0.
	for (int i0 = tid; i0 < G * HD; i0 += 2 * kCT) {
		const bool two = i0 + kCT < G * HD;
		const int i1 = two ? i0 + kCT : i0;
		const int g0 = i0 / HD, g1 = i1 / HD;
		const float *src = ws_acc + (size_t)kvh * ncta * G * HD;
		float v0 = 0.f, v1 = 0.f;
		for (int b = b_lo; b <= b_hi; b += kInFlight) {
			float p0[kInFlight], p1[kInFlight];
#pragma unroll
			for (int k = 0; k < kInFlight; ++k) {
				const int bk = min(b + k, b_hi);
				const int ba = wts[bk * G + g0] < 0.f ? b_hi : bk, bb = wts[bk * G + g1] < 0.f ? b_hi : bk;
				p0[k] = __ldcg(src + (size_t)ba * G * HD + i0);
				p1[k] = __ldcg(src + (size_t)bb * G * HD + i1);
			}
#pragma unroll
			for (int k = 0; k < kInFlight; ++k) {
				const int bk = min(b + k, b_hi);
				const bool in = b + k <= b_hi;
				const float w0 = in ? fmaxf(wts[bk * G + g0], 0.f) : 0.f, w1 = in ? fmaxf(wts[bk * G + g1], 0.f) : 0.f;
				v0 = fmaf(w0, p0[k], v0);
				v1 = fmaf(w1, p1[k], v1);
			}
		}
		const float o0 = v0 / wts[ncta * G + g0], o1 = v1 / wts[ncta * G + g1];
		att_out[(size_t)kvh * G * HD + i0] = o0;
		if (two) att_out[(size_t)kvh * G * HD + i1] = o1;
		if (ll_att) { // flagged hand-over: the wo staging step polls these
#pragma unroll
			for (int rp = 0; rp < kLLRep; ++rp) {
				ll_store(ll_att + (size_t)rp * n_q + (size_t)kvh * G * HD + i0, o0, wflag);
				if (two) ll_store(ll_att + (size_t)rp * n_q + (size_t)kvh * G * HD + i1, o1, wflag);
			}
		}
	}
}

// Distributed combine of one KV head (flagged hand-over): this CTA is participant `pi` of the head's `cnt` CTAs that own
// a chunk and produces outputs [pi, pi+1) * G*HD / cnt.  Called from the phase loop (in front of the wo staging step)
// with a recomputed split rather than from attention_consume: nothing of it may be live across the attention unit loop.
template <int HD, int G>
__device__ __forceinline__ void ll_combine(const unsigned long long *ll_ml, const unsigned long long *ll_acc, unsigned long long *ll_att,
                                        float *att_out, int kvh, int ncta, int c0, int nc, int nch, int pi, unsigned wflag, int tid, int n_q) {
	const int cnt = min(nc, nch);
	const int o_lo = (int)((unsigned)pi * (G * HD) / cnt), o_hi = (int)((unsigned)(pi + 1) * (G * HD) / cnt);
	constexpr int U = 4; // participants in flight per thread
	for (int o = o_lo + tid; o < o_hi; o += kCT) {
		const int g = o / HD;
		const unsigned long long *mlp = ll_ml + (((size_t)kvh * ncta + c0) * G + g) * 2; // + r * G * 2
		const unsigned long long *acp = ll_acc + ((size_t)kvh * ncta + c0) * G * HD + o; // + r * G * HD
		float M = -INFINITY;
		for (int r0 = 0; r0 < nc; r0 += U) {
			unsigned long long w[U];
#pragma unroll
			for (int j = 0; j < U; ++j) {
				const int r = r0 + j;
				if (r < nc && attn_has_unit(r, nc, nch)) w[j] = ll_load1(mlp + (size_t)r * G * 2);
			}
#pragma unroll
			for (int j = 0; j < U; ++j) {
				const int r = r0 + j;
				if (r < nc && attn_has_unit(r, nc, nch))
					M = fmaxf(M, ll_ok(w[j], wflag) ? ll_val(w[j]) : ll_wait1(mlp + (size_t)r * G * 2, wflag));
			}
		}
		float ls = 0.f, v = 0.f;
		for (int r0 = 0; r0 < nc; r0 += U) {
			ulonglong2 ml[U];
			unsigned long long ac[U];
#pragma unroll
			for (int j = 0; j < U; ++j) {
				const int r = r0 + j;
				if (r < nc && attn_has_unit(r, nc, nch)) {
					ml[j] = ll_load2(mlp + (size_t)r * G * 2);
					ac[j] = ll_load1(acp + (size_t)r * G * HD);
				}
			}
#pragma unroll
			for (int j = 0; j < U; ++j) {
				const int r = r0 + j;
				if (r < nc && attn_has_unit(r, nc, nch)) {
					const float m_r = ll_val(ml[j].x); // validated by the first pass
					const float l_r = ll_ok(ml[j].y, wflag) ? ll_val(ml[j].y) : ll_wait1(mlp + (size_t)r * G * 2 + 1, wflag);
					const float p_r = ll_ok(ac[j], wflag) ? ll_val(ac[j]) : ll_wait1(acp + (size_t)r * G * HD, wflag);
					const float wgt = m_r == -INFINITY ? 0.f : expf(m_r - M);
					ls = fmaf(wgt, l_r, ls);
					v = fmaf(wgt, p_r, v);
				}
			}
		}
		const float out = v / ls;
		att_out[(size_t)kvh * G * HD + o] = out;
#pragma unroll
		for (int rp = 0; rp < kLLRep; ++rp) ll_store(ll_att + (size_t)rp * n_q + (size_t)kvh * G * HD + o, out, wflag); // polled by the wo staging step
	}
}

// Split-K flash-decode over this CTA's (kv head, row chunk) units; arithmetic as src/infer.cpp:216-248 (fp32 dot over
// fp16 K, / sqrt(head_dim) folded into q, exp(s - max) / sum, fp32 mix over fp16 V).
// Warp w owns units w, w+W, ... of the CTA's contiguous unit range.  Per unit (one ring tile: C rows of K and of V):
//   A. partial dots: every lane holds 8 dims of every row of its row group for all G heads -> 8*G partial sums;
//   B. transposing butterfly over the LPR lanes of a row: each exchange halves the number of values a lane
//      carries, so 8*G values cost ~2 shuffles each instead of log2(LPR) — afterwards a lane owns the complete
//      scores of NF rows of ONE head (head = high bits of its lane index);
//   C. online softmax on the owned scores only (one head per lane: 1 max / 1 sum / NF+1 exps per lane);
//      probabilities and the rescale factors go through a per-warp scratch in shared memory;
//   D. acc[g][8 dims] += p[row][g] * V[row][dims] with the lane's 8 dims of each row of its row group.
// A warp that moves on to the next KV head parks its state in shared memory.  Warps, then CTAs, are merged with
// the usual (max, sum, acc) rescaling in a fixed order, so results are run-to-run identical.  The last CTA to
// publish a head's partial combines them (arrival counter).
template <int HD, int G, bool LL>
__device__ __forceinline__ void attention_consume(Ctx &c, const Args &a, const Layer &L, int layer) {
	constexpr int W = kConsumerWarps;
	constexpr int LPR = HD / 8, RPS = 32 / LPR; // lanes per row, rows per warp step (= row groups)
	constexpr int C = kStageBytes / (4 * HD);   // rows per tile
	constexpr int ST = C / RPS;                 // row steps per tile
	constexpr int V = ST * G;                   // partial sums per lane before the butterfly
	constexpr int NF = V / LPR;                 // complete scores per lane after it
	constexpr int SB = ST / NF;                 // lanes of a row that share a head (each owns NF of the ST rows)
	constexpr int WS = HD + 2;                  // per (slot, warp, head) record in shared memory: acc[HD], m, l
	constexpr int PS = C * G + G;               // per-warp scratch: p[C][G], corr[G]
	static_assert(ST == 8 && V % LPR == 0 && NF >= 1 && NF <= ST && SB * G == LPR, "unsupported head_dim / group combination");
	const int lane = c.lane, warp = c.warp, ncta = c.ncta;
	const int sub = lane % LPR, rgrp = lane / LPR;
	const int my_g = sub / SB, my_st0 = (sub % SB) * NF; // this lane owns scores of head my_g, row steps my_st0 .. +NF
	const int nch = (c.kv_len + C - 1) / C;
	const AttnSplit sp = attn_split(c.cta, ncta, a.n_kv_heads, nch);
	const int NU = sp.nu, kvh = sp.kvh;
	const uint32_t base = c.it;
	c.it = base + (uint32_t)NU;
	if (NU == 0) return;
	const float div = sqrtf((float)HD);
	float *wsm = c.vec;                                 // [W][G][WS] warp records
	float *wts = wsm + W * G * WS;                      // merge weights, then combine: [ncta][G] weights, Lsum[G]
	float *ps = wts + ncta * G + G + (size_t)warp * PS; // this warp's p / corr scratch
	if (lane < G) { // each warp initialises its own records, so a __syncwarp orders it against park()
		float *rec = wsm + (warp * G + lane) * WS;
		rec[HD] = -INFINITY, rec[HD + 1] = 0.f;
	}
	float2 acc[G][4]; // the output accumulators: this lane's 8 dims as 4 pairs
	// q / sqrt(HD) of the head group lives in shared memory and is re-read per unit (8 LDS.128): held in registers it
	// pushed the unit loop over the register ceiling, and the compiler's answer was to keep q in local memory instead
	float *qs = wsm + ((W * G * WS + ncta * G + G + W * PS + 3) & ~3); // 16-byte aligned
	const uint32_t qs_u32 = c.vec_u32 + (uint32_t)(qs - c.vec) * 4 + (uint32_t)sub * 32;
	float m = -INFINITY, l = 0.f; // running max / partial sum of head my_g (max identical in all lanes of the head)
	auto park = [&]() { // merge the warp's row groups and lanes, store the warp's state
#pragma unroll
		for (int o = LPR; o < 32; o <<= 1) {
			l += __shfl_xor_sync(kFull, l, o);
#pragma unroll
			for (int g = 0; g < G; ++g)
#pragma unroll
				for (int j = 0; j < 4; ++j) {
					acc[g][j].x += __shfl_xor_sync(kFull, acc[g][j].x, o);
					acc[g][j].y += __shfl_xor_sync(kFull, acc[g][j].y, o);
				}
		}
#pragma unroll
		for (int o = 1; o < SB; o <<= 1) l += __shfl_xor_sync(kFull, l, o);
		if (rgrp == 0) {
#pragma unroll
			for (int g = 0; g < G; ++g) {
				float *rec = wsm + (warp * G + g) * WS;
#pragma unroll
				for (int j = 0; j < 4; ++j) rec[sub * 8 + 2 * j] = acc[g][j].x, rec[sub * 8 + 2 * j + 1] = acc[g][j].y;
			}
			if (sub % SB == 0) {
				float *rec = wsm + (warp * G + my_g) * WS;
				rec[HD] = m, rec[HD + 1] = l;
			}
		}
	};
	__syncwarp();
	for (int e = c.tid * 4; e < G * HD; e += kCT * 4) {
		const float4 v = __ldcg(reinterpret_cast<const float4 *>(a.q + (size_t)kvh * G * HD + e));
		*reinterpret_cast<float4 *>(qs + e) = make_float4(v.x / div, v.y / div, v.z / div, v.w / div);
	}
#pragma unroll
	for (int g = 0; g < G; ++g)
#pragma unroll
		for (int j = 0; j < 4; ++j) acc[g][j] = make_float2(0.f, 0.f);
	cbar();
	// The unit loop needs every register it can get: the CTA-uniform context it does not use waits in shared memory
	// (the alternative is the compiler spilling inside the loop, and local memory is an L2 round trip here).
	volatile int *pk = sm_misc() + kMiscPark;
	if (c.tid == 0) {
		pk[0] = c.rot, pk[1] = (int)c.bar_gen, pk[2] = (int)c.xgen, sm_misc()[kMiscMmaTiles] = (int)c.mt;
		pk[3] = (int)(unsigned)((unsigned long long)c.prof), pk[4] = (int)(unsigned)((unsigned long long)c.prof >> 32);
	}
	for (int i = warp; i < NU; i += W) {
		const int t0 = (sp.k0 + i) * C, n = min(C, c.kv_len - t0);
		uint32_t slot;
		const unsigned char *tile = wait_tile(c, base + (uint32_t)i, &slot);
		const uint32_t ks = tile_u32(c, tile) + lane * 16, vs = ks + kStageBytes / 2; // row (st, rgrp) at + st * 512
		const size_t goff = ((size_t)kvh * a.max_seq_len + t0) * HD;
		const __half *kg = reinterpret_cast<const __half *>(L.kc) + goff;
		const __half *vg = reinterpret_cast<const __half *>(L.vc) + goff;
		// the row appended by this token and the re-rotated sink rows were written after the producer may have
		// fetched the tile: those come from global memory (L2), everything else from the ring
		const bool has_fresh = (c.kv_pos >= t0 && c.kv_pos < t0 + n) || t0 < c.kv_sink;
		// ---- A: partial dots, value index = g * ST + st ---------------------------------------------------------
		float d[V];
		float2 qr[G][4];
#pragma unroll
		for (int g = 0; g < G; ++g) {
			const float4 lo = lds128f(qs_u32 + g * HD * 4), hi = lds128f(qs_u32 + g * HD * 4 + 16);
			qr[g][0] = make_float2(lo.x, lo.y), qr[g][1] = make_float2(lo.z, lo.w);
			qr[g][2] = make_float2(hi.x, hi.y), qr[g][3] = make_float2(hi.z, hi.w);
		}
#pragma unroll
		for (int st = 0; st < ST; ++st) {
			const int row = st * RPS + rgrp, t = t0 + row;
			uint4 kv = lds128(ks + st * 512);
			if (has_fresh && row < n && (t == c.kv_pos || t < c.kv_sink))
				kv = __ldcg(reinterpret_cast<const uint4 *>(kg + (size_t)row * HD + sub * 8));
			const float2 k0 = WTraits<YALM_F16>::h2(kv.x), k1 = WTraits<YALM_F16>::h2(kv.y), k2 = WTraits<YALM_F16>::h2(kv.z),
			             k3 = WTraits<YALM_F16>::h2(kv.w);
#pragma unroll
			for (int g = 0; g < G; ++g) {
				float2 s2 = ffma2(qr[g][0], k0, make_float2(0.f, 0.f));
				s2 = ffma2(qr[g][1], k1, s2);
				s2 = ffma2(qr[g][2], k2, s2);
				s2 = ffma2(qr[g][3], k3, s2);
				d[g * ST + st] = s2.x + s2.y;
			}
		}
		// ---- B: transposing butterfly over the LPR lanes of a row ----------------------------------------------
		{
			int nv = V;
#pragma unroll
			for (int o = LPR / 2; o > 0; o >>= 1) {
				const bool up = (sub & o) != 0;
				const int half = nv / 2;
#pragma unroll
				for (int i2 = 0; i2 < V / 2; ++i2) {
					if (i2 < half) {
						const float send = up ? d[i2] : d[i2 + half];
						const float keep = up ? d[i2 + half] : d[i2];
						d[i2] = keep + __shfl_xor_sync(kFull, send, o);
					}
				}
				nv = half;
			}
		}
		// d[0..NF) = complete scores of head my_g, row steps my_st0 + i   (value index = g*ST + st, see above)
		// ---- C: online softmax for head my_g over the tile's rows ------------------------------------------------
		float mt = -INFINITY;
#pragma unroll
		for (int i2 = 0; i2 < NF; ++i2) {
			const int row = (my_st0 + i2) * RPS + rgrp;
			if (row >= n || (a.debug & 4)) d[i2] = -INFINITY;
			mt = fmaxf(mt, d[i2]);
		}
#pragma unroll
		for (int o = 1; o < SB; o <<= 1) mt = fmaxf(mt, __shfl_xor_sync(kFull, mt, o));
#pragma unroll
		for (int o = LPR; o < 32; o <<= 1) mt = fmaxf(mt, __shfl_xor_sync(kFull, mt, o));
		const float m_new = fmaxf(m, mt); // finite: every tile has at least one valid row
		const float corr = m == -INFINITY ? 0.f : __expf(m - m_new);
		m = m_new;
		l *= corr;
		__syncwarp(); // the previous unit's scratch reads are done
#pragma unroll
		for (int i2 = 0; i2 < NF; ++i2) {
			const float p = d[i2] == -INFINITY ? 0.f : __expf(d[i2] - m_new);
			l += p;
			ps[((my_st0 + i2) * RPS + rgrp) * G + my_g] = p;
		}
		if (rgrp == 0 && sub % SB == 0) ps[C * G + my_g] = corr;
		__syncwarp();
		// ---- D: rescale, then acc[g] += p[row][g] * V[row] --------------------------------------------------------
#pragma unroll
		for (int g = 0; g < G; ++g) {
			const float cg = ps[C * G + g];
#pragma unroll
			for (int j = 0; j < 4; ++j) acc[g][j].x *= cg, acc[g][j].y *= cg;
		}
#pragma unroll
		for (int st = 0; st < ST; ++st) {
			const int row = st * RPS + rgrp, t = t0 + row;
			if (row < n) {
				uint4 vv = lds128(vs + st * 512);
				if (has_fresh && (t == c.kv_pos || t < c.kv_sink))
					vv = __ldcg(reinterpret_cast<const uint4 *>(vg + (size_t)row * HD + sub * 8));
				const float2 v0 = WTraits<YALM_F16>::h2(vv.x), v1 = WTraits<YALM_F16>::h2(vv.y), v2 = WTraits<YALM_F16>::h2(vv.z),
				             v3 = WTraits<YALM_F16>::h2(vv.w);
#pragma unroll
				for (int g = 0; g < G; ++g) {
					const float p = ps[row * G + g];
					const float2 pp = make_float2(p, p);
					acc[g][0] = ffma2(pp, v0, acc[g][0]);
					acc[g][1] = ffma2(pp, v1, acc[g][1]);
					acc[g][2] = ffma2(pp, v2, acc[g][2]);
					acc[g][3] = ffma2(pp, v3, acc[g][3]);
				}
			}
		}
		release(c, slot);
	}
	if (warp < NU) park();
	cbar();
	c.rot = pk[0], c.bar_gen = (unsigned)pk[1], c.xgen = (unsigned)pk[2], c.mt = (unsigned)sm_misc()[kMiscMmaTiles];
	c.prof = reinterpret_cast<long long *>((unsigned long long)(unsigned)pk[3] | (unsigned long long)(unsigned)pk[4] << 32);
	stamp(c, layer, 1, 5);

	// ---- merge the warps, publish the CTA's partial, the last CTA of the head to arrive combines ------------
	const unsigned wflag = ll_flag(a.seq, layer * ppl_of(a) + 2); // flagged hand-over: this phase is ph = ppl * layer + 1
	constexpr bool LLC = LL && YALM_MEGA_LL_ATTN;
	float *ew = wts; // [W][G] merge weights (wts is not in use before the combine)
	if (c.tid < W * G) { // weight of warp w's record for head g, computed once
		const int w = c.tid / G, g = c.tid % G;
		float M = -INFINITY;
#pragma unroll
		for (int w2 = 0; w2 < W; ++w2) M = fmaxf(M, wsm[(w2 * G + g) * WS + HD]);
		const float mw = wsm[(w * G + g) * WS + HD];
		ew[c.tid] = mw == -INFINITY ? 0.f : expf(mw - M); // M is finite: warp 0 of a participating CTA owns a unit
		if (w == 0) {
			float ls = 0.f;
#pragma unroll
			for (int w2 = 0; w2 < W; ++w2) {
				const float m2 = wsm[(w2 * G + g) * WS + HD];
				if (m2 != -INFINITY) ls = fmaf(expf(m2 - M), wsm[(w2 * G + g) * WS + HD + 1], ls);
			}
			if (LLC) {
				ll_store2(a.ll_ml + (((size_t)kvh * ncta + c.cta) * G + g) * 2, M, ls, wflag);
			} else {
				float *pml = a.ws_ml + (((size_t)kvh * ncta + c.cta) * G + g) * 2;
				pml[0] = M, pml[1] = ls;
			}
		}
	}
	cbar();
	for (int i = c.tid; i < G * (HD / 2); i += kCT) { // two dims per thread
		const int g = i / (HD / 2), d = (i % (HD / 2)) * 2;
		float2 v = make_float2(0.f, 0.f);
#pragma unroll
		for (int w = 0; w < W; ++w) {
			const float e = ew[w * G + g];
			const float2 r = *reinterpret_cast<const float2 *>(wsm + (w * G + g) * WS + d);
			if (e != 0.f) v.x = fmaf(e, r.x, v.x), v.y = fmaf(e, r.y, v.y); // empty records hold no data
		}
		if (LLC) ll_store2(a.ll_acc + (((size_t)kvh * ncta + c.cta) * G + g) * HD + d, v.x, v.y, wflag);
		else *reinterpret_cast<float2 *>(a.ws_acc + (((size_t)kvh * ncta + c.cta) * G + g) * HD + d) = v;
	}
	stamp(c, layer, 1, 7);
	if (LLC) {
		// Distributed combine: the head's participating CTAs each own a slice of its G * HD outputs and poll the
		// partials (max, sum, acc) of all participants directly — no arrival counter, no barrier, no fence.  Summed in
		// CTA order with the same weights on every run: deterministic.
		stamp(c, layer, 1, 6);
		return;
	}
	cbar(); // the partial's stores happen-before the release below
	if (c.tid == 0) {
		const int cnt = min(sp.nc, nch); // CTAs of the head that own a chunk
		unsigned prev;
		asm volatile("atom.acq_rel.gpu.global.add.u32 %0, [%1], 1;" : "=r"(prev) : "l"(a.head_counters + kvh) : "memory");
		const bool last = prev + 1 == (unsigned)cnt;
		sm_misc()[0] = last ? 1 : 0;
		if (last) a.head_counters[kvh] = 0; // every participant has arrived: ready for the next layer
	}
	cbar();
	stamp(c, layer, 1, 6);
	if (sm_misc()[0]) // CTA-uniform
		combine_head<HD, G>(a.ws_acc, a.ws_ml, a.att_out, wts, wsm, kvh, ncta, nch, sp.c0, sp.c0 + sp.nc - 1,
		                    c.prof ? c.prof + ((size_t)layer * kProfPhases + 1) * kProfSlots + 4 : nullptr, LL ? a.ll_att : nullptr, wflag, a.q_dim);
}

// Mixture-of-experts router, run by every CTA on the normalised vector it has just staged (c.vec): the n_experts gate
// logits (one warp per expert row, weights from L2), then moe_gate (src/infer.cpp:100-132): top-k by repeated strict-'>'
// arg-max (lowest index wins ties), weights = softmax over the SELECTED experts with exponentials relative to the global
// maximum.  Same instruction sequence in every CTA -> the same bits everywhere; nothing crosses CTAs.  The record in
// sm_misc releases the producer warp, which has been waiting for the expert ids (Producer::expert).
template <int DT>
__device__ __forceinline__ void moe_router(Ctx &c, const Args &a, const Layer &L, int layer) {
	using T = WTraits<DT>;
	constexpr int EP = T::EP, CH = 32 * EP, Q = EP / 4;
	float *gl = sm_red(); // 16 floats: kMegaMaxExperts
	for (int e = c.warp; e < a.n_experts; e += kConsumerWarps) {
		const char *row = L.gate + (size_t)e * a.dim * T::BYTES;
		float2 acc = make_float2(0.f, 0.f);
		for (int c0 = 0; c0 < a.dim; c0 += CH) {
			if (c0 + c.lane * EP < a.dim) {
				const uint4 w = __ldcg(reinterpret_cast<const uint4 *>(row + (size_t)(c0 + c.lane * EP) * T::BYTES));
				float4 xv[Q];
#pragma unroll
				for (int q = 0; q < Q; ++q) xv[q] = lds128f(c.vec_u32 + (uint32_t)(c0 + q * 128) * 4 + c.lane * 16);
				acc = fma16x2<DT>(w, xv, acc);
			}
		}
		float v = warp_sum(acc.x + acc.y);
		if (DT == YALM_F8E4M3 && L.sgate) v *= L.sgate[e];
		if (c.lane == 0) gl[e] = v;
	}
	cbar();
	if (c.tid == 0) {
		volatile int *misc = sm_misc();
		float max_val = -FLT_MAX;
		for (int j = 0; j < a.n_experts; ++j)
			if (gl[j] > max_val) max_val = gl[j];
		unsigned mask = 0;
		float wsum = 0.f;
		for (int k = 0; k < a.n_active; ++k) {
			int best = -1;
			for (int j = 0; j < a.n_experts; ++j)
				if (!(mask >> j & 1u) && (best == -1 || gl[j] > gl[best])) best = j;
			misc[kMiscExpert + k] = best;
			wsum += expf(gl[best] - max_val);
			mask |= 1u << best;
		}
		for (int k = 0; k < a.n_active; ++k) {
			const int e = misc[kMiscExpert + k];
			const float w = expf(gl[e] - max_val) / wsum;
			misc[kMiscExpertW + k] = __float_as_int(w);
			if (c.cta == 0) a.active_e[k] = e, a.active_w[k] = w; // state taps (InferenceState::active_experts*)
		}
		if (c.cta == 0 && a.moe_logits)
			for (int j = 0; j < a.n_experts; ++j) a.moe_logits[j] = gl[j];
		__threadfence_block();
		misc[kMiscRouter] = layer + 1; // releases the producer warp
	}
	cbar();
}

__device__ __forceinline__ void derive_ring(int pos, int max_seq_len, int *kv_sink, int *kv_pos, int *kv_len) {
	// src/infer.cu:1081-1083
	*kv_sink = pos >= max_seq_len ? KV_SINKS : 0;
	*kv_pos = *kv_sink + (pos - *kv_sink) % (max_seq_len - *kv_sink);
	*kv_len = pos >= max_seq_len ? max_seq_len : pos + 1;
}

// ---- the kernel -----------------------------------------------------------------------------------------
template <int DT, int HD, int G, bool TP, int V>
__global__ void __launch_bounds__(kThreads, 1) decode_mega_kernel(const __grid_constant__ Args a) {
	using T = WTraits<DT>;
	constexpr int B = T::BYTES;
	Ctx c;
	c.S = a.stages;
	c.s_magic = (uint32_t)((0x100000000ull + (unsigned)a.stages - 1) / (unsigned)a.stages);
	c.vec = reinterpret_cast<float *>(sm_stages() + (size_t)c.S * kStageBytes);
	c.vec_u32 = sm_stages_u32() + (uint32_t)c.S * kStageBytes;
	c.cta = blockIdx.x, c.ncta = gridDim.x;
	c.tid = threadIdx.x, c.warp = threadIdx.x >> 5, c.lane = threadIdx.x & 31;
	c.it = 0, c.rot = 0, c.bar_gen = 0;
	c.xgen = 0, c.mt = 0;
	if (TP) c.xgen = *reinterpret_cast<const volatile unsigned *>(a.tp_peer[a.tp_rank] + kTpOffGen); // written by the previous launch
	c.prof = a.prof ? a.prof + (size_t)blockIdx.x * a.n_layers * kProfPhases * kProfSlots : nullptr;
#ifdef YALM_MEGA_TRACE
	c.wait_cycles = 0;
	c.trace = blockIdx.x == 0 ? a.trace : nullptr;
#endif

	int token = a.token, pos = a.pos;
	if (a.token < 0) token = a.ctl[CTL_TOKEN], pos = a.ctl[CTL_POS];
	derive_ring(pos, a.max_seq_len, &c.kv_sink, &c.kv_pos, &c.kv_len);
	if (threadIdx.x == 0) sm_misc()[kMiscToken] = token, sm_misc()[kMiscPos] = pos, sm_misc()[kMiscRouter] = 0; // token / pos: needed at phase 0 / at the very end only

	if (threadIdx.x == 0) {
		for (int i = 0; i < c.S; ++i) {
			mbar_init(sm_full0() + 8 * i, 1);
			mbar_init(sm_empty0() + 8 * i, 1);
			sm_rel()[i] = 0;
		}
		for (int i = 0; V == 1 && i < kHbBufs; ++i) {
			mbar_init(sm_hbfull0() + 8 * i, 1);
			mbar_init(sm_hbempty0() + 8 * i, kConsumerWarps);
		}
		for (int i = 0; V == 2 && i < kMmaSlots; ++i) {
			mbar_init(sm_mma_done0() + 8 * i, kConsumerWarps);
			mbar_init(sm_mma_free0() + 8 * i, 1);
		}
		asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
	}
	// RoPE table of this token: cos/sin(pos * theta^(-j/rotary_dim)) with the reference's fp32 formula
	// (src/infer.cpp:200-213; accurate powf/cosf/sinf: pos * freq reaches 3e4 rad at 32k context)
	if (threadIdx.x < HD / 2) {
		const int j = 2 * threadIdx.x;
		const float freq = j >= a.rotary_dim ? 0.f : 1.0f / powf(a.theta, (float)j / (float)a.rotary_dim);
		const float val = pos * freq;
		sm_rope()[threadIdx.x] = make_float2(cosf(val), sinf(val));
	}
	__syncthreads();

	if (c.warp >= kConsumerWarps) {
		// ================================ producers ================================
		// (the L2 look-ahead experiment does not know about expert-indirect rows: off for MoE)
		Producer p{c, (a.debug & 8) ? 0ull : policy_evict_first(), min(c.S, 32), (a.debug & 8) || a.n_experts > 0 ? 0 : abs(a.l2_ahead), a.l2_ahead > 0, (a.debug & 16) != 0};
		p.template run<V>(a, HD, B);
		return;
	}

	// ================================ consumers ================================
	Epi ep;
	const int n_phases = a.n_layers * ppl_of(a) + (a.mode != kHydrate ? 1 : 0);
	const bool moe = a.n_experts > 0;
	for (int ph = 0; ph < n_phases; ++ph) { // one loop body = one call site per phase function
		int l, k, sub;
		phase_of(ph, a, &l, &k, &sub);
		ep.sub = k == 4 ? sub : 0, ep.n_sub = k == 4 ? a.n_sub : 1, ep.ew = 1.0f, ep.gpe = 0;
		const Layer &L = a.layers[l];
		ep.L = &L, ep.layer = l;
		stamp(c, l, k, 0);
		const float *rms = k == 0 ? L.rms_att : k == 3 ? L.rms_ffn : k == 5 ? a.rms_final : nullptr;
		float4 wv[kNormCache];
		load_norm_weights(c, rms, a.dim, wv); // in flight across the barrier
		// flagged hand-over: only the QKV -> attention boundary keeps its (local) grid barrier, every other phase polls its
		// input vector in the staging step.  Tensor parallel: the vectors behind wo / w2 were finished by the exchange at
		// the end of those phases (gemv_consume) and travel the same way — no cross-GPU barrier at any phase boundary.
		constexpr bool LL = YALM_MEGA_LL != 0;
		const bool flagged = LL && ph > 0 && k != 1; // this phase polls its input instead of a barrier
		const unsigned rflag = ll_flag(a.seq, ph - sub); // what the phase before this one produced (w2 sub-phases: all of hb comes from the w1/w3 phase)
		if (ph > 0 && !flagged) grid_barrier(c, a.sync);
		stamp(c, l, k, 1);
		if (k == 1) { // attention: q comes from registers, nothing to stage
			stamp(c, l, k, 2);
			attention_consume<HD, G, YALM_MEGA_LL != 0>(c, a, L, l);
			stamp(c, l, k, 3);
			continue;
		}
		const float *src = a.x;
		int n = a.dim;
		switch (k) {
		case 0: ep.kind = kQKV, ep.scale0 = L.sq, ep.scale1 = L.sk, ep.scale2 = L.sv, ep.rows0 = a.q_dim, ep.rows1 = a.kv_dim; break;
		case 2: src = a.att_out, n = a.q_dim, ep.kind = kWO, ep.scale0 = L.so; break;
		case 3: ep.kind = kFFN, ep.scale0 = L.s1, ep.scale1 = L.s3, ep.gpe = moe ? a.shape[kFFN][3] / a.n_sub : 0; break;
		case 4:
			src = a.hb + (size_t)sub * a.hidden, n = a.hidden, ep.kind = kW2, ep.scale0 = L.s2; // MoE: the sub-th selected expert's slice of hb
			if (moe) {
				ep.ew = __int_as_float(sm_misc()[kMiscExpertW + sub]);
				if (L.s2) ep.scale0 = L.s2 + (size_t)sm_misc()[kMiscExpert + sub] * a.dim;
			}
			break;
		default: ep.kind = kLM, ep.scale0 = a.wcls_scale, ep.best_v = -FLT_MAX, ep.best_i = 0x7fffffff; break;
		}
		if (V == 1 && k == 3) hb_setup(a, c); // pipelined hb: ordered before the tile loop by the staging step's CTA barriers
		if (V == 1 && k == 4) {
			// pipelined hb: nothing to stage — the slices of hb arrive by bulk copy, each awaited where it is needed
		} else if (flagged) {
			if (YALM_MEGA_LL_ATTN && k == 2) { // distributed attention combine of this CTA's head slice (flags of phase ph - 1)
				const int C = kStageBytes / (4 * HD), nch = (c.kv_len + C - 1) / C;
				const AttnSplit sp = attn_split(c.cta, c.ncta, a.n_kv_heads, nch);
				if (sp.nu > 0)
					ll_combine<HD, G>(a.ll_ml, a.ll_acc, a.ll_att, a.att_out, sp.kvh, c.ncta, sp.c0, sp.nc, nch,
					                  nch >= sp.nc ? c.cta - sp.c0 : sp.k0, rflag, c.tid, a.q_dim);
			}
			const unsigned long long *lsrc = k == 2 ? a.ll_att : k == 3 ? a.ll_x : k == 4 ? a.ll_hb : a.ll_x + (size_t)kLLRep * a.dim;
			lsrc += (size_t)(c.cta % kLLRep) * (k == 4 ? a.hb_len : n); // this CTA's replica
			if (k == 4) lsrc += (size_t)sub * n;
			stage_vec_ll<DT, V>(c, a, lsrc, n, rms, rflag, wv, c.prof ? c.prof + ((size_t)l * kProfPhases + k) * kProfSlots : nullptr);
		} else {
			stage_vec<DT, V>(c, a, src, n, rms, ph == 0 ? sm_misc()[kMiscToken] : -1, wv);
		}
		if (moe && k == 3) moe_router<DT>(c, a, L, l); // on the vector just staged; releases the producer's w1/w3 and w2 tiles
		stamp(c, l, k, 2);
		if (k == 0 && c.kv_sink > 0 && c.cta == c.ncta - 1) {
			// rotate each sink key forward by one position, fp16 -> fp32 -> fp16 (src/infer.cpp:303-317)
			const int hp = HD >> 1, npairs = c.kv_sink * a.n_kv_heads * hp;
			for (int p = c.tid; p < npairs; p += kCT) {
				const int r = p / (a.n_kv_heads * hp), h = (p / hp) % a.n_kv_heads, i = (p % hp) * 2;
				__half2 *ptr = reinterpret_cast<__half2 *>(reinterpret_cast<__half *>(L.kc) + ((size_t)h * a.max_seq_len + r) * HD + i);
				float2 v = __half22float2(*ptr);
				rope_pair(v.x, v.y, i, 1, a.theta, a.rotary_dim);
				*ptr = __floats2half2_rn(v.x, v.y);
			}
		}
		gemv_phase<DT, TP, V>(c, a, ep, ph);
		stamp(c, l, k, 3);
	}
	if (a.mode == kGreedy) {
		// greedy arg-max, first maximum wins (src/sampler.cpp:27-38): lanes -> warps -> CTA -> grid
		float bv = ep.best_v;
		int bi = ep.best_i;
#pragma unroll
		for (int o = 16; o > 0; o >>= 1) {
			const float ov = __shfl_xor_sync(kFull, bv, o);
			const int oi = __shfl_xor_sync(kFull, bi, o);
			if (ov > bv || (ov == bv && oi < bi)) bv = ov, bi = oi;
		}
		if (c.lane == 0) sm_red()[c.warp] = bv, sm_misc()[4 + c.warp] = bi;
		cbar();
		if (c.tid == 0) {
			bv = sm_red()[0], bi = sm_misc()[4];
			for (int t = 1; t < kConsumerWarps; ++t)
				if (sm_red()[t] > bv || (sm_red()[t] == bv && sm_misc()[4 + t] < bi)) bv = sm_red()[t], bi = sm_misc()[4 + t];
			if (!TP) a.amax_val[c.cta] = bv, a.amax_idx[c.cta] = bi;
			sm_red()[0] = bv, sm_misc()[4] = bi;
		}
		if (TP) { // the CTA's candidate goes to every rank: each reduces all world * ncta of them to the same token
			cbar();
			if (c.tid < a.tp_world) {
				char *peer = a.tp_peer[c.tid];
				reinterpret_cast<float *>(peer + kTpOffAmaxV)[a.tp_rank * kTpMaxGrid + c.cta] = sm_red()[0];
				reinterpret_cast<int *>(peer + kTpOffAmaxI)[a.tp_rank * kTpMaxGrid + c.cta] = sm_misc()[4];
			}
			tp_barrier(c, a); // also: every rank's logits rows have landed here
		} else {
			grid_barrier(c, a.sync);
		}
		if (c.cta == 0 && c.warp == 0) {
			bv = -FLT_MAX, bi = 0x7fffffff;
			if (TP) {
				const char *mine = a.tp_peer[a.tp_rank];
				for (int b = c.lane; b < c.ncta * a.tp_world; b += 32) {
					const int at = b / c.ncta * kTpMaxGrid + b % c.ncta;
					const float v = __ldcg(reinterpret_cast<const float *>(mine + kTpOffAmaxV) + at);
					const int i = __ldcg(reinterpret_cast<const int *>(mine + kTpOffAmaxI) + at);
					if (v > bv || (v == bv && i < bi)) bv = v, bi = i;
				}
			} else
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
				const int pos1 = sm_misc()[kMiscPos] + 1;
				a.ctl[CTL_POS] = pos1;
				int ks, kp, kl;
				derive_ring(pos1, a.max_seq_len, &ks, &kp, &kl);
				a.ctl[CTL_SINK] = ks, a.ctl[CTL_KVPOS] = kp, a.ctl[CTL_KVLEN] = kl;
			}
		}
	}
	// every launch ends on a cross-GPU barrier (greedy: the one above): all ranks' logits rows have landed in this
	// rank's region, and no rank starts pushing the next token's partials while another still reads this token's
	if (TP && a.mode != kGreedy) tp_barrier(c, a);
	// ---- leave the barrier words zeroed for the next launch ------------------------------------------
	cbar();
	if (c.tid == 0) {
		__threadfence();
		const unsigned prev = atomicAdd(a.sync + 1, 1u);
		const bool last = prev + 1 == (unsigned)c.ncta;
		if (last) {
			a.sync[0] = 0;
			a.sync[1] = 0;
			if (TP) *reinterpret_cast<unsigned *>(a.tp_peer[a.tp_rank] + kTpOffGen) = c.xgen;
			__threadfence();
		}
		sm_misc()[0] = last ? 1 : 0;
	}
	if (V == 1) { // the last CTA out zeroes the hb slice counters for the next launch
		cbar();
		if (sm_misc()[0])
			for (int i = c.tid; i < a.n_layers * kHbMaxSlices; i += kCT) a.hb_cnt[i] = 0;
	}
}

// (head_dim, q heads per kv head) combinations compiled; others use the per-op kernel path.  V: kernel variant (gemv_phase);
// the optional variants exist for one GPU only.
template <int DT, int V>
inline const void *kernel_for(int head_dim, int g, bool tp) {
	if (V != 0 && tp) return nullptr;
#define YALM_MEGA_CASE(HD, GG)                                                                                \
	if (head_dim == HD && g == GG) {                                                                          \
		if constexpr (V == 0) {                                                                               \
			if (tp) return (const void *)decode_mega_kernel<DT, HD, GG, true, 0>;                             \
		}                                                                                                     \
		return (const void *)decode_mega_kernel<DT, HD, GG, false, V>;                                        \
	}
	YALM_MEGA_CASE(128, 4) YALM_MEGA_CASE(128, 8) YALM_MEGA_CASE(64, 4) YALM_MEGA_CASE(32, 4) YALM_MEGA_CASE(32, 2)
#undef YALM_MEGA_CASE
	return nullptr;
}

} // namespace mega
} // namespace yalm
