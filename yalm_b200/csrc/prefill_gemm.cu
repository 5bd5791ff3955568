// Dense fp16 GEMM on the 5th-generation tensor cores for the batched prefill / perplexity path.
//
// The reference has no batched pass at all: it hydrates the KV cache and scores a text one token at a time, every token a
// full sweep over the weights (src/main.cpp:91-97, 174-184).  With M >= 128 positions in flight the projections ARE
// GEMMs, C[M, N] = A[M, K] * W[N, K]^T — W exactly as the .yalm file stores it ((out, in) row-major = K-major "B") — and
// this is the one place in the project where tensor cores are the right tool.
//
// One CTA per 128 x 128 output tile (128 x 256 accumulator columns for the fused w1 | w3 tile), 192 threads:
//   warp 0      TMA producer: cp.async.bulk.tensor.2d of a 128-row x 64-column fp16 box of A and of W per pipeline stage
//               (SWIZZLE_128B tensor maps; rows / columns past the edge are zero-filled by the TMA unit), mbarrier
//               complete_tx completion
//   warp 1      one elected thread issues tcgen05.mma.cta_group::1.kind::f16 (M 128, N 128 | 256, K 16) straight out of
//               shared memory (K-major SWIZZLE_128B descriptors), fp32 accumulator in TMEM; tcgen05.commit hands each
//               stage back to the producer and finally the accumulator to the epilogue.  Also owns the TMEM allocation.
//   warps 2-5   epilogue: tcgen05.ld (32 lanes x 32 columns per instruction) -> registers -> fused tail -> global:
//               plain fp32 store (q | k | v, logits), fp32 residual add in place (wo, w2: x += ...), or
//               act(w1 x) * (w3 x) -> fp16 (the A operand of the w2 GEMM).
// Activations enter as fp16 (rmsnorm / attention / GLU outputs are rounded once, like the KV cache); sums are fp32.
#include "common.cuh"

#include <cuda.h> // CUtensorMap + enums only: cuTensorMapEncodeTiled is resolved at run time (cudaGetDriverEntryPoint)

#include <mutex>

#include "gemv.cuh"

namespace yalm {
namespace prefill {

constexpr int BM = 128, BN = 128, BK = 64; // BK fp16 = 128 bytes = one SWIZZLE_128B row
constexpr int kGemmThreads = 192;
constexpr int kTileBytes = BM * BK * 2;    // 16 KB: one 128-row operand box

enum Tail { kStoreF32 = 0, kAddF32 = 1, kGluF16 = 2 };

struct GemmArgs {
	int M, N, K; // valid rows of A, rows of W (per matrix), inner dimension
	int tail, act;
	float *c;    // kStoreF32 / kAddF32: (M, ldc) fp32
	__half *h;   // kGluF16: (M, ldc) fp16
	int ldc;
};

// ---- PTX ---------------------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t s32(const void *p) { return static_cast<uint32_t>(__cvta_generic_to_shared(p)); }
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
	asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
	asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
	uint32_t ok = 0;
	const long long t0 = clock64();
	while (!ok) {
		asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
		             : "=r"(ok)
		             : "r"(bar), "r"(parity)
		             : "memory");
		if (!ok && clock64() - t0 > 4000000000ll) __trap(); // ~2 s: a bug, never a hang
	}
}
__device__ __forceinline__ void tma_load_2d(uint32_t dst, const CUtensorMap *map, int c0, int c1, uint32_t bar) {
	asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];" ::"r"(dst),
	             "l"(map), "r"(c0), "r"(c1), "r"(bar)
	             : "memory");
}
__device__ __forceinline__ void tmem_alloc(uint32_t slot, uint32_t cols) {
	asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(slot), "r"(cols) : "memory");
	asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t addr, uint32_t cols) {
	asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(addr), "r"(cols) : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
// D[tmem] (+)= A[smem] * B[smem]^T, one 128 x N x 16 fp16 step
__device__ __forceinline__ void umma_f16(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
	asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(tmem_d),
	             "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
	             : "memory");
}
// arrive on `bar` once every tcgen05.mma issued so far by this thread has completed (implies fence::before_thread_sync)
__device__ __forceinline__ void umma_commit(uint32_t bar) {
	asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
// 32 lanes x 32 consecutive fp32 columns of the accumulator -> 32 registers per thread (thread = TMEM lane = tile row)
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, float (&v)[32]) {
	uint32_t r[32];
	asm volatile("tcgen05.ld.sync.aligned.32x32b.x32.b32 "
	             "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
	             "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
	             : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]), "=r"(r[9]),
	               "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]), "=r"(r[17]), "=r"(r[18]),
	               "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]),
	               "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
	             : "r"(taddr)
	             : "memory");
	asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
	for (int i = 0; i < 32; ++i) v[i] = __uint_as_float(r[i]);
}

// Shared-memory matrix descriptor of a K-major operand tile in the canonical SWIZZLE_128B layout the TMA box lands in:
// rows of 128 bytes, 8-row groups 1024 bytes apart (stride byte offset), version 1 (sm_100), layout type 2 (128B swizzle).
// Fields (cute/arch/mma_sm100_desc.hpp: SmemDescriptor): [0,14) address >> 4, [16,30) leading byte offset >> 4 (1: unused
// for swizzled K-major), [32,46) stride byte offset >> 4, [46,48) version, [61,64) layout type.
__device__ __forceinline__ uint64_t smem_desc_sw128(uint32_t addr) {
	return (uint64_t)((addr & 0x3FFFF) >> 4) | (1ull << 16) | ((uint64_t)(1024 >> 4) << 32) | (1ull << 46) | (2ull << 61);
}
// Instruction descriptor (InstrDescriptor): fp32 accumulator [4,6) = 1, A / B format fp16 = 0, both K-major, N >> 3 at
// [17,23), M >> 4 at [24,29).
__host__ __device__ constexpr uint32_t instr_desc_f16(int M, int N) {
	return (1u << 4) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}

// NB = 1: one W matrix, accumulator 128 x 128.  NB = 2: w1 | w3 tiles side by side, accumulator 128 x 256, GLU tail.
template <int NB>
__global__ void __launch_bounds__(kGemmThreads, 1)
gemm_f16_kernel(const __grid_constant__ CUtensorMap ta, const __grid_constant__ CUtensorMap tb, const __grid_constant__ CUtensorMap tb2, const GemmArgs g) {
	constexpr int kStage = kTileBytes * (1 + NB);
	constexpr int kStages = NB == 1 ? 6 : 4;
	constexpr int NACC = BN * NB; // accumulator columns (fp32) = UMMA N
	extern __shared__ uint8_t smem_raw[];
	const uint32_t base = (s32(smem_raw) + 1023u) & ~1023u; // the swizzle pattern repeats every 1024 bytes
	const uint32_t bars = base + kStages * kStage;           // full[kStages] | empty[kStages] | accumulator ready | TMEM address
	const uint32_t full0 = bars, empty0 = bars + 8 * kStages, acc_bar = bars + 16 * kStages, tslot = acc_bar + 8;
	const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
	const int m0 = blockIdx.x * BM, n0 = blockIdx.y * BN;
	const int nkb = (g.K + BK - 1) / BK;

	if (threadIdx.x == 0) {
		for (int i = 0; i < kStages; ++i) mbar_init(full0 + 8 * i, 1), mbar_init(empty0 + 8 * i, 1);
		mbar_init(acc_bar, 1);
		asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
	}
	if (warp == 1) tmem_alloc(tslot, NACC);
	tc_fence_before();
	__syncthreads();
	tc_fence_after();
	uint32_t tmem;
	asm volatile("ld.shared.u32 %0, [%1];" : "=r"(tmem) : "r"(tslot));

	if (warp == 0) {
		if (lane == 0) { // ---- TMA producer ----
			for (int kb = 0; kb < nkb; ++kb) {
				const int s = kb % kStages, r = kb / kStages;
				if (r > 0) mbar_wait(empty0 + 8 * s, (uint32_t)(r - 1) & 1u);
				const uint32_t dst = base + s * kStage, bar = full0 + 8 * s;
				mbar_expect_tx(bar, kStage);
				tma_load_2d(dst, &ta, kb * BK, m0, bar);
				tma_load_2d(dst + kTileBytes, &tb, kb * BK, n0, bar);
				if (NB == 2) tma_load_2d(dst + 2 * kTileBytes, &tb2, kb * BK, n0, bar);
			}
		}
	} else if (warp == 1) {
		if (lane == 0) { // ---- MMA issuer ----
			constexpr uint32_t idesc = instr_desc_f16(BM, NACC);
			for (int kb = 0; kb < nkb; ++kb) {
				const int s = kb % kStages, r = kb / kStages;
				mbar_wait(full0 + 8 * s, (uint32_t)r & 1u);
				tc_fence_after();
				const uint32_t a = base + s * kStage, b = a + kTileBytes;
#pragma unroll
				for (int k = 0; k < BK / 16; ++k) // 16 fp16 = 32 bytes further along K, inside the 128-byte swizzle row
					umma_f16(tmem, smem_desc_sw128(a + 32 * k), smem_desc_sw128(b + 32 * k), idesc, (kb | k) != 0);
				umma_commit(empty0 + 8 * s); // the stage is free once these MMAs have read it
			}
			umma_commit(acc_bar); // the accumulator is complete
		}
	} else { // ---- epilogue: warp w reads TMEM lanes 32 * (w % 4) .. + 31 (the hardware's lane quarter of a warp) ----
		mbar_wait(acc_bar, 0);
		tc_fence_after();
		const int q = warp & 3, row = m0 + q * 32 + lane;
		const uint32_t trow = tmem + ((uint32_t)(q * 32) << 16);
#pragma unroll 1
		for (int c0 = 0; c0 < BN; c0 += 32) {
			float v[32];
			tmem_ld32(trow + c0, v);
			if (NB == 2) {
				float u[32];
				tmem_ld32(trow + BN + c0, u);
#pragma unroll
				for (int i = 0; i < 32; ++i) v[i] = (g.act == YALM_ACT_GELU ? act_gelu(v[i]) : act_silu(v[i])) * u[i];
			}
			if (row < g.M) {
				const int col = n0 + c0;
				if (g.tail == kGluF16) {
					__half *dst = g.h + (size_t)row * g.ldc + col;
#pragma unroll
					for (int i = 0; i < 32; i += 8) {
						if (col + i + 8 <= g.N) {
							uint4 pk;
							__half2 h0 = __floats2half2_rn(v[i], v[i + 1]), h1 = __floats2half2_rn(v[i + 2], v[i + 3]);
							__half2 h2 = __floats2half2_rn(v[i + 4], v[i + 5]), h3 = __floats2half2_rn(v[i + 6], v[i + 7]);
							pk.x = *reinterpret_cast<uint32_t *>(&h0), pk.y = *reinterpret_cast<uint32_t *>(&h1);
							pk.z = *reinterpret_cast<uint32_t *>(&h2), pk.w = *reinterpret_cast<uint32_t *>(&h3);
							*reinterpret_cast<uint4 *>(dst + i) = pk;
						} else {
							for (int j = i; j < i + 8 && col + j < g.N; ++j) dst[j] = __float2half_rn(v[j]);
						}
					}
				} else {
					float *dst = g.c + (size_t)row * g.ldc + col;
#pragma unroll
					for (int i = 0; i < 32; i += 4) {
						if (col + i + 4 <= g.N) {
							float4 o = make_float4(v[i], v[i + 1], v[i + 2], v[i + 3]);
							if (g.tail == kAddF32) { // x += W v (src/infer.cpp:335-337), residual stream stays fp32
								const float4 x = *reinterpret_cast<const float4 *>(dst + i);
								o.x += x.x, o.y += x.y, o.z += x.z, o.w += x.w;
							}
							*reinterpret_cast<float4 *>(dst + i) = o;
						} else {
							for (int j = i; j < i + 4 && col + j < g.N; ++j) dst[j] = g.tail == kAddF32 ? dst[j] + v[j] : v[j];
						}
					}
				}
			}
		}
		tc_fence_before();
	}
	__syncthreads();
	if (warp == 1) {
		tc_fence_after();
		tmem_dealloc(tmem, NACC);
	}
}

// ---- host side ------------------------------------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *,
                                  const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn encode_fn() {
	static EncodeTiledFn fn = nullptr;
	static std::once_flag once;
	std::call_once(once, [] {
		void *p = nullptr;
		cudaDriverEntryPointQueryResult q;
		if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess && q == cudaDriverEntryPointSuccess)
			fn = reinterpret_cast<EncodeTiledFn>(p);
	});
	return fn;
}

// (rows, cols) row-major fp16 matrix -> tensor map with a 128-row x 64-column box, 128-byte swizzle, zero fill past the edges
static int make_map(CUtensorMap *map, const void *ptr, int rows, int cols) {
	EncodeTiledFn fn = encode_fn();
	if (!fn) return YALM_EINVAL;
	const cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
	const cuuint64_t strides[1] = {(cuuint64_t)cols * 2};
	const cuuint32_t box[2] = {BK, BM}, estr[2] = {1, 1};
	const CUresult r = fn(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, const_cast<void *>(ptr), dims, strides, box, estr,
	                      CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
	                      CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
	return r == CUDA_SUCCESS ? 0 : YALM_EINVAL;
}

template <int NB>
static int launch(const CUtensorMap &ta, const CUtensorMap &tb, const CUtensorMap &tb2, const GemmArgs &g, cudaStream_t st) {
	constexpr int kStage = kTileBytes * (1 + NB), kStages = NB == 1 ? 6 : 4;
	const size_t smem = (size_t)kStages * kStage + 1024 + 256;
	static bool configured = false;
	if (!configured) {
		YALM_TRY(cudaFuncSetAttribute(gemm_f16_kernel<NB>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
		configured = true;
	}
	const dim3 grid((g.M + BM - 1) / BM, (g.N + BN - 1) / BN);
	gemm_f16_kernel<NB><<<grid, kGemmThreads, smem, st>>>(ta, tb, tb2, g);
	return (int)cudaGetLastError();
}

} // namespace prefill
} // namespace yalm

using namespace yalm;
using namespace yalm::prefill;

// c (M, ldc) fp32 (=|+=) a (M, K) fp16 * w (N, K)^T fp16.  a, w, c: device pointers, 16-byte aligned rows (K % 8 == 0).
extern "C" int yalm_gemm_f16(float *c, int ldc, const void *a, const void *w, int M, int N, int K, int accumulate, yalm_stream_t s) {
	if (!c || !a || !w || M < 1 || N < 1 || K < 8 || K % 8 || ldc < N || ldc % 4) return YALM_EINVAL;
	CUtensorMap ta, tb;
	int rc = make_map(&ta, a, M, K);
	if (!rc) rc = make_map(&tb, w, N, K);
	if (rc) return rc;
	GemmArgs g{M, N, K, accumulate ? kAddF32 : kStoreF32, 0, c, nullptr, ldc};
	return launch<1>(ta, tb, tb, g, (cudaStream_t)s);
}

// h (M, ldh) fp16 = act(a w1^T) * (a w3^T): the fused w1 / w3 projection + GLU of the FFN (src/infer.cpp:360-375).
extern "C" int yalm_gemm_glu_f16(void *h, int ldh, const void *a, const void *w1, const void *w3, int M, int N, int K, int act,
                                 yalm_stream_t s) {
	if (!h || !a || !w1 || !w3 || M < 1 || N < 1 || K < 8 || K % 8 || ldh < N || ldh % 8) return YALM_EINVAL;
	CUtensorMap ta, tb, tb2;
	int rc = make_map(&ta, a, M, K);
	if (!rc) rc = make_map(&tb, w1, N, K);
	if (!rc) rc = make_map(&tb2, w3, N, K);
	if (rc) return rc;
	GemmArgs g{M, N, K, kGluF16, act, nullptr, static_cast<__half *>(h), ldh};
	return launch<2>(ta, tb, tb2, g, (cudaStream_t)s);
}

// Host-pointer test entry: c (M, N) fp32 = a (M, K) fp16 * w (N, K)^T fp16 [+ c]; w3 != NULL: h (M, N) fp16 = GLU.
extern "C" int yalm_gemm_host(float *c, void *h, const void *a, const void *w, const void *w3, int M, int N, int K, int accumulate,
                              int act) {
	int n = 0;
	if (cudaGetDeviceCount(&n) != cudaSuccess || n == 0) return YALM_ENOGPU;
	if ((!c && !h) || !a || !w || M < 1 || N < 1 || K < 8) return YALM_EINVAL;
	void *da = nullptr, *dw = nullptr, *dw3 = nullptr, *dc = nullptr;
	const size_t ab = (size_t)M * K * 2, wb = (size_t)N * K * 2, cb = (size_t)M * N * (w3 ? 2 : 4);
	YALM_TRY(cudaMalloc(&da, ab));
	YALM_TRY(cudaMalloc(&dw, wb));
	YALM_TRY(cudaMalloc(&dc, cb));
	YALM_TRY(cudaMemcpy(da, a, ab, cudaMemcpyHostToDevice));
	YALM_TRY(cudaMemcpy(dw, w, wb, cudaMemcpyHostToDevice));
	int rc;
	if (w3) {
		YALM_TRY(cudaMalloc(&dw3, wb));
		YALM_TRY(cudaMemcpy(dw3, w3, wb, cudaMemcpyHostToDevice));
		rc = yalm_gemm_glu_f16(dc, N, da, dw, dw3, M, N, K, act, nullptr);
	} else {
		YALM_TRY(cudaMemcpy(dc, c, cb, cudaMemcpyHostToDevice));
		rc = yalm_gemm_f16(static_cast<float *>(dc), N, da, dw, M, N, K, accumulate, nullptr);
	}
	if (!rc) rc = (int)cudaDeviceSynchronize();
	if (!rc) rc = (int)cudaMemcpy(w3 ? h : (void *)c, dc, cb, cudaMemcpyDeviceToHost);
	cudaFree(da), cudaFree(dw), cudaFree(dw3), cudaFree(dc);
	return rc;
}
