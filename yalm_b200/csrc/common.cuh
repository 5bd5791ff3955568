// Shared device helpers for the sm_100a decode kernels: streaming 128-bit loads, warp/block
// reductions, weight-type traits (fp32 / fp16 / fp8-e4m3 with fused dequant).
#pragma once

#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>

#include "../../include/yalm_b200.h"

// Reference error convention for the memory helpers (src/infer.cu:13-31): message + abort().
#define YALM_CUDA_ABORT(x)                                                                                    \
	do {                                                                                                      \
		cudaError_t err__ = (x);                                                                              \
		if (err__ != cudaSuccess) {                                                                           \
			fprintf(stderr, "CUDA error in %s at %s:%d: %s (%s=%d)\n", __FUNCTION__, __FILE__, __LINE__,       \
			        cudaGetErrorString(err__), cudaGetErrorName(err__), (int)err__);                          \
			abort();                                                                                          \
		}                                                                                                     \
	} while (0)

// Op-level convention: return the error to the caller.
#define YALM_TRY(x)                                                                                           \
	do {                                                                                                      \
		cudaError_t err__ = (x);                                                                              \
		if (err__ != cudaSuccess) return (int)err__;                                                          \
	} while (0)

namespace yalm {

constexpr int kWarp = 32;
constexpr unsigned kFull = 0xffffffffu;
constexpr int KV_SINKS = 2; // src/model.h:12

// ctl block slots (include/yalm_b200.h)
enum { CTL_TOKEN = 0, CTL_POS = 1, CTL_SINK = 2, CTL_KVPOS = 3, CTL_KVLEN = 4, CTL_STEP = 5 };

// 128-bit read-only streaming load: weights are read exactly once per token and are ~100x larger
// than L2, so they must not displace activations / KV in L1.
__device__ __forceinline__ uint4 ldg_stream(const void *p) {
	uint4 r;
	asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
	             : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
	             : "l"(p));
	return r;
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
	for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(kFull, v, o);
	return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
	for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(kFull, v, o));
	return v;
}

// Block-wide sum, result broadcast to every thread. `red` = 33 floats of shared memory.
// Deterministic: fixed tree over warps.
__device__ __forceinline__ float block_sum(float v, float *red) {
	const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
	v = warp_sum(v);
	__syncthreads(); // protect `red` from a previous use
	if (lane == 0) red[wid] = v;
	__syncthreads();
	if (wid == 0) {
		float t = lane < nw ? red[lane] : 0.f;
		t = warp_sum(t);
		if (lane == 0) red[32] = t;
	}
	__syncthreads();
	return red[32];
}
__device__ __forceinline__ float block_max(float v, float *red) {
	const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
	v = warp_max(v);
	__syncthreads();
	if (lane == 0) red[wid] = v;
	__syncthreads();
	if (wid == 0) {
		float t = lane < nw ? red[lane] : -INFINITY;
		t = warp_max(t);
		if (lane == 0) red[32] = t;
	}
	__syncthreads();
	return red[32];
}

// ---- weight element traits ---------------------------------------------------------------------
// EP = elements per 16-byte load.  fma16(): acc += dot(the EP elements of `w`, xv[0..EP)).
template <int DT>
struct WTraits;

template <>
struct WTraits<YALM_F32> {
	static constexpr int EP = 4;
	static constexpr int BYTES = 4;
	__device__ static __forceinline__ float fma16(uint4 w, const float4 *xv, float acc) {
		acc = fmaf(__uint_as_float(w.x), xv[0].x, acc);
		acc = fmaf(__uint_as_float(w.y), xv[0].y, acc);
		acc = fmaf(__uint_as_float(w.z), xv[0].z, acc);
		acc = fmaf(__uint_as_float(w.w), xv[0].w, acc);
		return acc;
	}
	__device__ static __forceinline__ float to_float(const void *p, size_t i) {
		return static_cast<const float *>(p)[i];
	}
};

template <>
struct WTraits<YALM_F16> {
	static constexpr int EP = 8;
	static constexpr int BYTES = 2;
	__device__ static __forceinline__ float2 h2(uint32_t u) {
		return __half22float2(*reinterpret_cast<const __half2 *>(&u));
	}
	__device__ static __forceinline__ float fma16(uint4 w, const float4 *xv, float acc) {
		float2 a = h2(w.x), b = h2(w.y), c = h2(w.z), d = h2(w.w);
		acc = fmaf(a.x, xv[0].x, acc);
		acc = fmaf(a.y, xv[0].y, acc);
		acc = fmaf(b.x, xv[0].z, acc);
		acc = fmaf(b.y, xv[0].w, acc);
		acc = fmaf(c.x, xv[1].x, acc);
		acc = fmaf(c.y, xv[1].y, acc);
		acc = fmaf(d.x, xv[1].z, acc);
		acc = fmaf(d.y, xv[1].w, acc);
		return acc;
	}
	__device__ static __forceinline__ float to_float(const void *p, size_t i) {
		return __half2float(static_cast<const __half *>(p)[i]);
	}
};

template <>
struct WTraits<YALM_F8E4M3> {
	static constexpr int EP = 16;
	static constexpr int BYTES = 1;
	// two e4m3 codes (low 16 bits) -> two floats, exact (hardware cvt e4m3x2 -> f16x2 -> f32)
	__device__ static __forceinline__ float2 e2(uint16_t u) {
		uint32_t h;
		asm("cvt.rn.f16x2.e4m3x2 %0, %1;" : "=r"(h) : "h"(u));
		return __half22float2(*reinterpret_cast<const __half2 *>(&h));
	}
	__device__ static __forceinline__ float word(uint32_t w, float4 xv, float acc) {
		float2 a = e2((uint16_t)(w & 0xffffu)), b = e2((uint16_t)(w >> 16));
		acc = fmaf(a.x, xv.x, acc);
		acc = fmaf(a.y, xv.y, acc);
		acc = fmaf(b.x, xv.z, acc);
		acc = fmaf(b.y, xv.w, acc);
		return acc;
	}
	__device__ static __forceinline__ float fma16(uint4 w, const float4 *xv, float acc) {
		acc = word(w.x, xv[0], acc);
		acc = word(w.y, xv[1], acc);
		acc = word(w.z, xv[2], acc);
		acc = word(w.w, xv[3], acc);
		return acc;
	}
	__device__ static __forceinline__ float to_float(const void *p, size_t i) {
		uint16_t u = static_cast<const uint8_t *>(p)[i];
		return e2(u).x;
	}
};

// ---- the activation vector in shared memory ----------------------------------------------------
// A lane's 16-byte weight load covers EP consecutive elements, so it needs EP consecutive x values =
// EP/4 LDS.128.  Stored naively those LDS.128 are 2-way (fp16) / 4-way (fp8) bank conflicted (lane
// stride 32/64 bytes).  We store x permuted inside each 32*EP-element chunk so that quarter q of
// every lane's slice is contiguous across lanes:  slot(c, q, lane, j) = c*32*EP + q*128 + lane*4 + j
// holds x[c*32*EP + lane*EP + q*4 + j].  Every LDS.128 is then conflict-free.
template <int EP>
__device__ __forceinline__ int xs_slot(int e) {
	constexpr int CH = 32 * EP;
	const int c = e / CH, r = e % CH;
	const int lane = r / EP, q = (r % EP) >> 2, j = r & 3;
	return c * CH + q * 128 + lane * 4 + j;
}
template <int EP>
__host__ __device__ constexpr int xs_padded(int n) {
	return (n + 32 * EP - 1) / (32 * EP) * (32 * EP);
}

} // namespace yalm
