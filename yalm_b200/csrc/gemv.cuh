// Bandwidth-bound GEMV building blocks: one warp streams R weight rows with 128-bit loads
// (R*U loads in flight per lane), multiplies against the CTA's shared activation vector, and
// all-reduces with shuffles.  fp32 accumulate everywhere, like the reference (src/infer.cu:216-232).
#pragma once

#include "common.cuh"

namespace yalm {

// Stage x (n floats, n % 4 == 0) into the permuted shared layout (see xs_slot).
template <int EP>
__device__ __forceinline__ void stage_x_plain(float *xs, const float *__restrict__ x, int n) {
	for (int e = threadIdx.x * 4; e < n; e += blockDim.x * 4) {
		float4 v = *reinterpret_cast<const float4 *>(x + e);
		*reinterpret_cast<float4 *>(xs + xs_slot<EP>(e)) = v;
	}
}

// xs = rmsnorm(x) * weight, the reference's formula and association order
// (src/infer.cu:526-539 / src/infer.cpp:134-144): scale = 1/sqrt(sum(x^2)/n + eps); o = x*scale*w.
// Every CTA recomputes it from L2 (n*4 bytes) instead of a separate single-CTA launch.
template <int EP>
__device__ __forceinline__ void stage_x_rmsnorm(float *xs, const float *__restrict__ x,
                                                const float *__restrict__ weight, int n, float eps, float *red) {
	float ss = 0.f;
	for (int e = threadIdx.x * 4; e < n; e += blockDim.x * 4) {
		float4 v = *reinterpret_cast<const float4 *>(x + e);
		ss += v.x * v.x + v.y * v.y + v.z * v.z + v.w * v.w;
	}
	ss = block_sum(ss, red);
	const float scale = 1.0f / sqrtf(ss / n + eps);
	for (int e = threadIdx.x * 4; e < n; e += blockDim.x * 4) {
		float4 v = *reinterpret_cast<const float4 *>(x + e);
		float4 w = *reinterpret_cast<const float4 *>(weight + e);
		float4 o = make_float4(v.x * scale * w.x, v.y * scale * w.y, v.z * scale * w.z, v.w * scale * w.w);
		*reinterpret_cast<float4 *>(xs + xs_slot<EP>(e)) = o;
	}
}

// acc[r] = dot(row r, x) for R rows given as byte pointers; result valid in every lane.
// Preconditions: n % EP == 0, rows 16-byte aligned.
template <int DT, int R, int U>
__device__ __forceinline__ void warp_rows_dot(const char *const (&rows)[R], const float *xs, int n,
                                              float (&acc)[R]) {
	using T = WTraits<DT>;
	constexpr int EP = T::EP, CH = 32 * EP, Q = EP / 4;
	const int lane = threadIdx.x & 31;
	const int nfull = n / CH;
	const size_t lane_off = (size_t)lane * 16;
#pragma unroll
	for (int r = 0; r < R; ++r) acc[r] = 0.f;

	int c = 0;
	for (; c + U <= nfull; c += U) {
		uint4 w[U][R];
#pragma unroll
		for (int u = 0; u < U; ++u)
#pragma unroll
			for (int r = 0; r < R; ++r) w[u][r] = ldg_stream(rows[r] + (size_t)(c + u) * 512 + lane_off);
		// Scheduling fence: without it the compiler sinks the loads between the FMAs to save registers
		// and only 4-6 of the R*U loads are ever in flight.  All loads issue first, math after.
#pragma unroll
		for (int u = 0; u < U; ++u)
#pragma unroll
			for (int r = 0; r < R; ++r)
				asm volatile("" : "+r"(w[u][r].x), "+r"(w[u][r].y), "+r"(w[u][r].z), "+r"(w[u][r].w));
#pragma unroll
		for (int u = 0; u < U; ++u) {
			float4 xv[Q];
#pragma unroll
			for (int q = 0; q < Q; ++q)
				xv[q] = *reinterpret_cast<const float4 *>(xs + (c + u) * CH + q * 128 + lane * 4);
#pragma unroll
			for (int r = 0; r < R; ++r) acc[r] = T::fma16(w[u][r], xv, acc[r]);
		}
	}
	// leftover full chunks (< U of them) and the ragged last chunk
	for (; c * CH < n; ++c) {
		if (c * CH + lane * EP < n) {
			uint4 w[R];
#pragma unroll
			for (int r = 0; r < R; ++r) w[r] = ldg_stream(rows[r] + (size_t)c * 512 + lane_off);
			float4 xv[Q];
#pragma unroll
			for (int q = 0; q < Q; ++q) xv[q] = *reinterpret_cast<const float4 *>(xs + c * CH + q * 128 + lane * 4);
#pragma unroll
			for (int r = 0; r < R; ++r) acc[r] = T::fma16(w[r], xv, acc[r]);
		}
	}
#pragma unroll
	for (int r = 0; r < R; ++r) acc[r] = warp_sum(acc[r]);
}

// activations: src/infer.cu:586-596 (same constants as the CPU path, src/infer.cpp:187-193)
__device__ __forceinline__ float act_silu(float x) {
	return x / (1.0f + expf(-x));
}
__device__ __forceinline__ float act_gelu(float x) {
	float x3 = x * x * x;
	return 0.5f * x * (1.0f + tanhf(0.797885f * (x + 0.044715f * x3)));
}
__device__ __forceinline__ float clipf(float x, float v) { // src/infer.cpp:195-197
	return x < -v ? -v : (x > v ? v : x);
}

// Interleaved-pair RoPE of one pair whose even element sits at index j_head inside its head
// (src/infer.cpp:200-213, src/infer.cu:541-556).  Accurate powf/cosf/sinf on purpose: pos*freq
// reaches 3e4 rad at 32k context and the fast intrinsics lose the parity tolerance there.
__device__ __forceinline__ void rope_pair(float &v0, float &v1, int j_head, int pos, float theta, int rotary_dim) {
	float freq = j_head >= rotary_dim ? 0.f : 1.0f / powf(theta, (float)j_head / (float)rotary_dim);
	float val = pos * freq;
	float fcr = cosf(val), fci = sinf(val);
	float a = v0 * fcr - v1 * fci;
	float b = v0 * fci + v1 * fcr;
	v0 = a;
	v1 = b;
}

} // namespace yalm
