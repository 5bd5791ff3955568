// Host side of the whole-token persistent decode kernel (planning, launch) and the streaming probe.
// The kernel itself is in decode_mega.cuh, instantiated per weight type in decode_mega_{f32,f16,f8}.cu.
#include "decode_mega.cuh"
#include "launch.h"

#include <cuda.h> // CUtensorMap + enums only: cuTensorMapEncodeTiled is resolved at run time (cudaGetDriverEntryPoint)

#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>

namespace yalm {
namespace mega {

// Pure streaming probe: the same ring and bulk copies over a flat buffer, consumers only wait and release.
// Its GB/s is the ceiling the megakernel can reach on this part (yalm_bench_op("stream")).
__global__ void __launch_bounds__(kThreads, 1) stream_probe_kernel(const char *buf, size_t bytes, int stages, unsigned *sink,
                                                                      int copy_bytes, int ahead) {
	extern __shared__ __align__(128) unsigned char smem_raw[];
	uint64_t *bars = reinterpret_cast<uint64_t *>(smem_raw + (size_t)stages * kStageBytes);
	const uint32_t full0 = smem_u32(bars), empty0 = full0 + 8 * kMaxStages, st0 = smem_u32(smem_raw);
	const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
	if (threadIdx.x == 0) {
		for (int i = 0; i < stages; ++i) {
			mbar_init(full0 + 8 * i, 1);
			mbar_init(empty0 + 8 * i, kConsumerWarps);
		}
		asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
	}
	__syncthreads();
	const size_t ntiles = bytes / kStageBytes;
	if (warp >= kConsumerWarps) {
		if (lane != 0 || warp > kConsumerWarps) return;
		const uint64_t pol = policy_evict_first();
		uint32_t it = 0;
		for (size_t t = blockIdx.x; t < ntiles; t += gridDim.x, ++it) {
			const uint32_t s = it % stages, par = ((it / stages) & 1u) ^ 1u;
			if (ahead > 0 && t + (size_t)ahead * gridDim.x < ntiles) // experiment: L2 prefetch `ahead` tiles in front of the ring
				bulk_prefetch_l2(buf + (t + (size_t)ahead * gridDim.x) * kStageBytes, kStageBytes);
			mbar_wait(empty0 + 8 * s, par);
			mbar_expect_tx(full0 + 8 * s, kStageBytes);
			for (int o = 0; o < kStageBytes; o += copy_bytes) // one stage as kStageBytes / copy_bytes bulk copies
				bulk_g2s(st0 + s * kStageBytes + o, buf + t * kStageBytes + o, copy_bytes, full0 + 8 * s, pol);
		}
		return;
	}
	uint32_t it = 0;
	unsigned acc = 0;
	for (size_t t = blockIdx.x; t < ntiles; t += gridDim.x, ++it) {
		const uint32_t s = it % stages, par = (it / stages) & 1u;
		mbar_wait(full0 + 8 * s, par);
		acc += *reinterpret_cast<const unsigned *>(smem_raw + (size_t)s * kStageBytes + threadIdx.x * 16);
		__syncwarp();
		if (lane == 0) mbar_arrive(empty0 + 8 * s);
	}
	if (acc == 0x12345678u) *sink = acc;
}

// ---- failure record -------------------------------------------------------------------------------------------
unsigned *fault_record() {
	static unsigned *rec = nullptr;
	if (!rec) {
		if (cudaHostAlloc((void **)&rec, 8 * sizeof(unsigned), cudaHostAllocMapped | cudaHostAllocPortable) != cudaSuccess) {
			cudaGetLastError();
			rec = nullptr;
		} else
			memset(rec, 0, 8 * sizeof(unsigned));
	}
	return rec;
}
long long fault_timeout_cycles() {
	const char *e = getenv("YALM_MEGA_TIMEOUT_S");
	return e && atof(e) > 0 ? (long long)(atof(e) * 2.0e9) : 0;
}
const char *describe_fault() {
	static char buf[256];
	const unsigned *r = fault_record();
	if (!r || !r[0]) return "";
	static const char *names[] = {"?", "ring stage (mbarrier)", "producer: stage release counter", "consumer: stage release counter",
	                              "flagged hand-over poll", "grid barrier", "cross-GPU barrier", "tensor-parallel mailbox", "MoE router record",
	                              "hand-over sample", "hb slice counter"};
	snprintf(buf, sizeof(buf), "decode_mega_kernel timed out in: %s (code %u) CTA %u thread %u values 0x%x %u", names[r[0] < 11 ? r[0] : 0], r[0],
	         r[1], r[2], r[3], r[4]);
	return buf;
}

// ---- host side --------------------------------------------------------------------------------------------
int make_weight_map(void *map128, const void *dev_ptr, int rows, int K, int box_rows) {
	typedef CUresult (*EncodeTiledFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *,
	                                  const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle,
	                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
	static EncodeTiledFn fn = [] {
		void *p = nullptr;
		cudaDriverEntryPointQueryResult q;
		if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess) p = nullptr;
		return reinterpret_cast<EncodeTiledFn>(p);
	}();
	static_assert(sizeof(CUtensorMap) == 128, "tensor map records are 128 bytes");
	if (!fn || K % 64 || box_rows < 1 || box_rows > 256) return YALM_EINVAL;
	// (column within a 64-column block, row, column block): a box = 64 x box_rows x 4 lands as [block][row][128 bytes]
	// (tried: (column, column block, row) boxes, which the copy engine walks row by row — 512 contiguous bytes at a time — and
	// which land as [row][block][128 bytes]: the same streaming rate, 4-way bank conflicts in ldmatrix)
	const cuuint64_t dims[3] = {64, (cuuint64_t)rows, (cuuint64_t)(K / 64)};
	const cuuint64_t strides[2] = {(cuuint64_t)K * 2, 128};
	const cuuint32_t box[3] = {64, (cuuint32_t)box_rows, kMmaKCols / 64}, estr[3] = {1, 1, 1};
	const CUresult r = fn(reinterpret_cast<CUtensorMap *>(map128), CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 3, const_cast<void *>(dev_ptr), dims, strides,
	                      box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
	                      CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
	return r == CUDA_SUCCESS ? 0 : YALM_EINVAL;
}

static inline int ep_of(int dtype) {
	return dtype == YALM_F32 ? 4 : dtype == YALM_F16 ? 8 : dtype == YALM_F8E4M3 ? 16 : 0;
}
static inline int bytes_of(int dtype) {
	return dtype == YALM_F32 ? 4 : dtype == YALM_F16 ? 2 : 1;
}
static inline int padded(int n, int ep) {
	const int ch = 32 * ep;
	return (n + ch - 1) / ch * ch;
}

size_t attn_ws_floats(int grid, int n_heads, int head_dim) {
	return (size_t)grid * n_heads * (head_dim + 2);
}

const void *kernel_f32(int head_dim, int g, bool tp);
const void *kernel_f16(int head_dim, int g, bool tp);
const void *kernel_f8(int head_dim, int g, bool tp);
const void *kernel_f32_pipe(int head_dim, int g, bool tp);
const void *kernel_f16_pipe(int head_dim, int g, bool tp);
const void *kernel_f8_pipe(int head_dim, int g, bool tp);
const void *kernel_f16_mma(int head_dim, int g, bool tp);

// variant: 0 default, 1 pipelined hb, 2 tensor-core GEMV (decode_mega.cuh: gemv_phase); null if not compiled
static const void *kernel_of(int dtype, int head_dim, int g, bool tp, int variant) {
	switch (dtype) {
	case YALM_F32: return variant == 0 ? kernel_f32(head_dim, g, tp) : variant == 1 ? kernel_f32_pipe(head_dim, g, tp) : nullptr;
	case YALM_F16:
		return variant == 0 ? kernel_f16(head_dim, g, tp) : variant == 1 ? kernel_f16_pipe(head_dim, g, tp) : kernel_f16_mma(head_dim, g, tp);
	case YALM_F8E4M3: return variant == 0 ? kernel_f8(head_dim, g, tp) : variant == 1 ? kernel_f8_pipe(head_dim, g, tp) : nullptr;
	}
	return nullptr;
}

// Tile shapes of the five GEMV kinds (qkv, wo, w1/w3, w2, lm head) for one rank's dimensions: rows per tile, columns per
// tile, K splits, row groups.  Pure arithmetic (no device): plan() and the CPU tests (yalm_mega_tile_plan) share it.
static bool plan_shapes(int dtype, int dim, int q_dim, int kv_dim, int hidden, int vocab, int n_sub, int (*shape)[4]) {
	const int ep = ep_of(dtype), B = bytes_of(dtype);
	if (!ep) return false;
	Shape s[kNumKinds];
	if (!tile_shape(q_dim, kv_dim, dim, B, ep, false, true, &s[kQKV])) return false;
	if (!tile_shape(dim, 0, q_dim, B, ep, false, false, &s[kWO])) return false;
	if (!tile_shape(hidden, 0, dim, B, ep, true, true, &s[kFFN])) return false;
	if (!tile_shape(dim, 0, hidden, B, ep, false, false, &s[kW2])) return false;
	if (!tile_shape(vocab, 0, dim, B, ep, false, false, &s[kLM])) return false;
	// MoE: the w1/w3 phase covers the n_sub selected experts (n_sub x hidden rows of hb); w2 runs once per expert
	s[kQKV].G = (q_dim + 2 * kv_dim) / s[kQKV].R, s[kWO].G = dim / s[kWO].R, s[kFFN].G = n_sub * (hidden / (s[kFFN].R / 2));
	s[kW2].G = dim / s[kW2].R, s[kLM].G = vocab / s[kLM].R;
	for (int k = 0; k < kNumKinds; ++k) shape[k][0] = s[k].R, shape[k][1] = s[k].KT, shape[k][2] = s[k].nk, shape[k][3] = s[k].G;
	return true;
}

Plan plan(int dtype, int dim, int q_dim, int kv_dim, int head_dim, int hidden, int vocab, int n_heads, int n_kv_heads,
          int n_experts, int n_active, int max_seq_len, int n_layers, bool tp, int full_vocab) {
	Plan p;
	const int n_sub = n_experts > 0 ? n_active : 1;
	if (n_experts < 0 || n_experts > kMegaMaxExperts || (n_experts > 0 && (n_active < 1 || n_active > kMegaMaxActive || n_active > n_experts))) return p;
	if (n_layers < 1 || n_layers * (4 + n_sub) + 2 >= (1 << kFlagPhaseBits)) return p; // phase field of the hand-over flags
	if (tp && (dim > kTpCapFloats || full_vocab > kTpMaxVocab)) return p;
	p.tp = tp;
	const int ep = ep_of(dtype);
	if (!ep) return p;
	if (n_kv_heads <= 0 || n_heads % n_kv_heads) return p;
	const int G = n_heads / n_kv_heads;
	if (!kernel_of(dtype, head_dim, G, tp, 0)) return p;
	if (dim % ep || q_dim % ep || hidden % ep || dim % 4 || q_dim % 4 || hidden % 4) return p;
	if (!plan_shapes(dtype, dim, q_dim, kv_dim, hidden, vocab, n_sub, p.shape)) return p;
	const DeviceInfo &di = device_info();
	int coop = 0, dev = 0;
	cudaGetDevice(&dev);
	cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, dev);
	if (!coop || di.sms < 1) return p;
	p.grid = di.sms;
	if (const char *e = getenv("YALM_MEGA_GRID")) // experiment: fewer CTAs than SMs (row blocks that fill whole 16-row tiles)
		if (atoi(e) >= 8 && atoi(e) < p.grid) p.grid = atoi(e);
	if (tp && p.grid > kTpMaxGrid) return p;
	if (n_kv_heads > p.grid) return p; // a CTA's attention units must span at most kHeadSlots KV heads
	// the attention unit split runs in 32-bit arithmetic: (units) x (grid + 1) must stay below 2^31
	const long long units = (long long)n_kv_heads * (max_seq_len / (kStageBytes / (4 * head_dim)) + 1);
	if (head_dim <= 0 || units * (p.grid + 1) >= (1ll << 31)) return p;
	// attention: warp records, combine weights, per-warp probability scratch (C rows x G + G)
	const int vf_attn = kHeadSlots * kConsumerWarps * G * (head_dim + 2) + p.grid * G + G +
	                    kConsumerWarps * (kStageBytes / (4 * head_dim) * G + G) + G * head_dim + 32;
	// header (rope table, scratch, release counters, mbarriers) | ring | vec
	auto stages_for = [&](int vf) {
		const long room = (long)di.max_smem_optin - ((long)((std::max(vf, vf_attn) + 31) / 32 * 32) * 4 + kSmemHeader);
		return (int)std::min<long>(kMaxStages, room / kStageBytes);
	};
	int vf = std::max(padded(dim, ep), std::max(padded(q_dim, ep), padded(hidden, ep)));
	// Tensor-core GEMV (decode_mega.cuh "tensor-core GEMV"): fp16 dense models, every K a multiple of 256 columns.
	// YALM_MEGA_MMA=1 / 0 force it on (where the shape allows) / off.
	{
		const char *e = getenv("YALM_MEGA_MMA");
		const bool can = dtype == YALM_F16 && !tp && n_experts == 0 && dim % kMmaKCols == 0 && q_dim % kMmaKCols == 0 &&
		                 hidden % kMmaKCols == 0 && q_dim % 2 == 0 && kv_dim % 2 == 0 && dim % 2 == 0 && vocab % 2 == 0;
		if (can && (e ? atoi(e) != 0 : false)) {
			p.mma = 1;
			auto box_rows = [&](int rows, int unit, int cap) { return mma_box_rows(rows, unit, cap, p.grid); };
			p.mma_mt[0] = box_rows(q_dim, 2, 16), p.mma_mt[1] = box_rows(kv_dim, 2, 16), p.mma_mt[2] = box_rows(dim, 2, 16);
			p.mma_mt[3] = box_rows(hidden, 1, 8), p.mma_mt[4] = box_rows(dim, 2, 16), p.mma_mt[5] = box_rows(vocab, 2, 16);
			vf += kMmaSlots * kConsumerWarps * 16; // the tiles' per-warp partial sums live behind the input vector
		}
	}
	// Pipelined w1/w3 -> w2 hand-over (decode_mega.cuh "pipelined hb"; dense models whose w2 is K-split): hb never sits in
	// shared memory as a whole — kHbBufs slice buffers behind the other phases' input vector replace it, and the ring gets the
	// difference (70B-class: 112 KB -> 64 KB, six more stages; measured +6.5 % tok/s on a 16-layer slice).  It is switched
	// on where it buys at least two ring stages (Mistral-7B: one stage, measured at parity: off); YALM_MEGA_W2_PIPE=1 / 0
	// force it on (where the shape allows) / off.
	{
		const char *e = getenv("YALM_MEGA_W2_PIPE");
		const int nk = p.shape[kW2][2], groups = (p.shape[kW2][3] + p.grid - 1) / std::max(1, p.grid); // row groups of w2 per CTA
		const int hb_off = std::max(padded(dim, ep), padded(q_dim, ep)), vf_pipe = hb_off + kHbBufs * p.shape[kW2][1];
		const bool can = !p.mma && !tp && n_experts == 0 && nk >= 2 && nk <= kHbMaxSlices && p.shape[kW2][0] <= 2 &&
		                 (groups + kConsumerWarps - 1) / kConsumerWarps <= kHbMaxGroups && stages_for(vf_pipe) >= 2;
		const bool want = e ? atoi(e) != 0 : stages_for(vf_pipe) >= stages_for(vf) + 2;
		if (can && want) p.w2_pipe = 1, p.hb_slice = p.shape[kW2][1], p.hb_off = hb_off, vf = vf_pipe;
	}
	p.kernel = kernel_of(dtype, head_dim, G, tp, p.mma ? 2 : p.w2_pipe ? 1 : 0);
	if (!p.kernel) return p;
	p.vec_floats = (std::max(vf, vf_attn) + 31) / 32 * 32;
	const size_t fixed = (size_t)p.vec_floats * 4 + kSmemHeader;
	p.stages = stages_for(vf);
	if (p.stages < 2) return p;
	p.smem = (size_t)p.stages * kStageBytes + fixed;
	p.dtype = dtype;
	if (cudaFuncSetAttribute(p.kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)p.smem) != cudaSuccess) {
		cudaGetLastError();
		return p;
	}
	int occ = 0;
	if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, p.kernel, kThreads, p.smem) != cudaSuccess || occ < 1) {
		cudaGetLastError();
		return p;
	}
	p.ok = true;
	return p;
}

int launch(const Plan &p, const Args &a, cudaStream_t st) {
	if (!p.ok) return YALM_EINVAL;
	Args args = a;
	args.stages = p.stages, args.vec_floats = p.vec_floats;
	args.w2_pipe = p.w2_pipe, args.hb_slice = p.hb_slice, args.hb_off = p.hb_off;
	args.mma = p.mma;
	memcpy(args.mma_mt, p.mma_mt, sizeof(args.mma_mt));
	if (p.mma && !a.tmaps) return YALM_EINVAL;
	if (p.w2_pipe && (!a.hbp || !a.hb_cnt)) return YALM_EINVAL;
	memcpy(args.shape, p.shape, sizeof(args.shape));
	// Timing-experiment knobs.  Read once; with YALM_MEGA_TUNE=1 (tools/mega_sweep.py) they are re-read at every launch.
	struct Knobs {
		int debug, max_stages, ahead;
		static int env(const char *name) {
			const char *v = getenv(name);
			return v ? atoi(v) : 0;
		}
		static Knobs read() { return {env("YALM_MEGA_DEBUG"), env("YALM_MEGA_STAGES"), env("YALM_MEGA_L2_AHEAD")}; }
	};
	static const bool tune = Knobs::env("YALM_MEGA_TUNE") != 0;
	static const Knobs fixed = Knobs::read();
	const Knobs kn = tune ? Knobs::read() : fixed;
	const int max_stages = kn.max_stages;
	args.debug = kn.debug;
	args.l2_ahead = kn.ahead;
	if (max_stages >= 2 && max_stages < args.stages) args.stages = max_stages;
	void *params[] = {&args};
	return (int)cudaLaunchCooperativeKernel(p.kernel, dim3(p.grid), dim3(kThreads), params, p.smem, st);
}

} // namespace mega
} // namespace yalm

// Host-only planning arithmetic of the whole-token kernel, for tests without a device.
extern "C" int yalm_mega_tile_plan(int dtype, int dim, int q_dim, int kv_dim, int hidden, int vocab, int *out20) {
	if (!out20) return YALM_EINVAL;
	int shape[5][4];
	if (!yalm::mega::plan_shapes(dtype, dim, q_dim, kv_dim, hidden, vocab, 1, shape)) return YALM_EINVAL;
	for (int k = 0; k < 5; ++k)
		for (int j = 0; j < 4; ++j) out20[4 * k + j] = shape[k][j];
	return 0;
}
extern "C" int yalm_mega_mma_plan(int rows, int unit, int cap, int grid, int cta, int *out4) {
	if (!out4 || rows < 1 || unit < 1 || cap < unit || rows % unit || grid < 1 || cta < 0 || cta >= grid) return YALM_EINVAL;
	int row0, nrows;
	yalm::mega::mma_block(cta, grid, rows, unit, &row0, &nrows);
	const int mt = yalm::mega::mma_box_rows(rows, unit, cap, grid);
	out4[0] = mt, out4[1] = row0, out4[2] = nrows, out4[3] = (nrows + mt - 1) / mt;
	return 0;
}
extern "C" int yalm_mega_hb_share(int groups, int rows_per_group, int slice_elems, int grid, int cta, int rotation, int slice) {
	if (groups < 1 || rows_per_group < 1 || slice_elems < rows_per_group || slice_elems % rows_per_group || grid < 1 || cta < 0 || cta >= grid ||
	    rotation < 0 || rotation >= grid || slice < 0)
		return YALM_EINVAL;
	const int g_first = (cta - rotation + grid) % grid, NL = g_first < groups ? (groups - g_first + grid - 1) / grid : 0;
	return yalm::mega::hb_share(g_first, NL, grid, slice_elems / rows_per_group, rows_per_group, slice);
}
extern "C" int yalm_mega_attn_split(int cta, int ncta, int n_kv_heads, int n_chunks, int *out5) {
	if (!out5 || ncta < 1 || cta < 0 || cta >= ncta || n_kv_heads < 1 || n_kv_heads > ncta || n_chunks < 1) return YALM_EINVAL;
	const yalm::mega::AttnSplit sp = yalm::mega::attn_split(cta, ncta, n_kv_heads, n_chunks);
	out5[0] = sp.kvh, out5[1] = sp.c0, out5[2] = sp.nc, out5[3] = sp.k0, out5[4] = sp.nu;
	return 0;
}

// Streaming ceiling of the ring (GB/s = bytes / usec / 1e3): `iters` timed launches over a `bytes` buffer.
extern "C" int yalm_stream_probe(size_t bytes, int iters, double *usec) {
	using namespace yalm;
	using namespace yalm::mega;
	if (!usec || iters < 1 || bytes < (size_t)kStageBytes) return YALM_EINVAL;
	int copy_bytes = getenv("YALM_PROBE_COPY") ? atoi(getenv("YALM_PROBE_COPY")) : kStageBytes;
	if (copy_bytes < 16 || copy_bytes > kStageBytes || kStageBytes % copy_bytes) copy_bytes = kStageBytes;
	const int ahead = getenv("YALM_PROBE_AHEAD") ? atoi(getenv("YALM_PROBE_AHEAD")) : 0;
	const DeviceInfo &di = device_info();
	const int stages = (int)std::min<long>(kMaxStages, ((long)di.max_smem_optin - 256) / kStageBytes);
	const size_t smem = (size_t)stages * kStageBytes + 2 * kMaxStages * 8;
	YALM_TRY(cudaFuncSetAttribute(stream_probe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
	char *buf = nullptr;
	unsigned *sink = nullptr;
	YALM_TRY(cudaMalloc(&buf, bytes));
	YALM_TRY(cudaMalloc(&sink, 4));
	YALM_TRY(cudaMemset(buf, 1, bytes));
	cudaEvent_t e0, e1;
	cudaEventCreate(&e0);
	cudaEventCreate(&e1);
	for (int i = 0; i < 2; ++i) stream_probe_kernel<<<di.sms, kThreads, smem>>>(buf, bytes, stages, sink, copy_bytes, ahead);
	cudaEventRecord(e0);
	for (int i = 0; i < iters; ++i) stream_probe_kernel<<<di.sms, kThreads, smem>>>(buf, bytes, stages, sink, copy_bytes, ahead);
	cudaEventRecord(e1);
	cudaError_t err = cudaDeviceSynchronize();
	float ms = 0.f;
	cudaEventElapsedTime(&ms, e0, e1);
	cudaEventDestroy(e0);
	cudaEventDestroy(e1);
	cudaFree(buf);
	cudaFree(sink);
	*usec = 1000.0 * ms / iters;
	return (int)err;
}
