// Sampling on the device: the reference copies vocab fp32 logits to the host every token and scans
// them there (src/sampler.cpp:11-65).  Here the next token is chosen where the logits are and 4 bytes travel.
//   sample_kernel       Sampler::sample (src/sampler.cpp:40-65): p_i = exp((l_i - max) / T) / sum, the first i whose
//                       running sum reaches r wins, vocab - 1 if none does.  r comes from the caller (tests) or from a
//                       counter-based generator keyed by (seed, step) — the reference draws std::rand(), which is not
//                       reproducible across runs either (seeded with the wall clock, src/main.cpp:50).
//   token_probs_kernel  Sampler::sample_prob (src/sampler.cpp:11-25) for a batch of logit rows: what perplexity mode
//                       needs per position (src/main.cpp:174-184), one float back per row instead of a vocab row.
// temperature == 0 is the greedy path (yalm_argmax_advance).
#include "common.cuh"

#include <cfloat>

namespace yalm {

__device__ __forceinline__ void derive_ring_ctl(int *ctl, int pos, int max_seq_len) { // src/infer.cu:1081-1083
	const int sink = pos >= max_seq_len ? KV_SINKS : 0;
	ctl[CTL_SINK] = sink;
	ctl[CTL_KVPOS] = sink + (pos - sink) % (max_seq_len - sink);
	ctl[CTL_KVLEN] = pos >= max_seq_len ? max_seq_len : pos + 1;
}

// splitmix64 -> uniform float in [0, 1)
__host__ __device__ inline float uniform_of(unsigned long long seed, unsigned long long counter) {
	unsigned long long z = seed + 0x9e3779b97f4a7c15ull * (counter + 1);
	z = (z ^ (z >> 30)) * 0xbf58476d1ce4e5b9ull;
	z = (z ^ (z >> 27)) * 0x94d049bb133111ebull;
	z ^= z >> 31;
	return (float)(z >> 40) * (1.0f / 16777216.0f);
}

constexpr int kSampleThreads = 1024;

// One CTA.  r_in < 0: r = uniform_of(seed, ctl ? ctl.step : counter).  ctl == null: only *token_out is written.
__global__ void __launch_bounds__(kSampleThreads) sample_kernel(const float *__restrict__ logits, int vocab, float temperature,
                                                                float r_in, unsigned long long seed, int *ctl, int max_seq_len,
                                                                int *tokens_out, int *token_out, unsigned long long counter) {
	__shared__ float red[33];
	__shared__ float wtot[32];
	__shared__ int first;
	const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
	float m = -FLT_MAX;
	for (int i = tid; i < vocab; i += kSampleThreads) m = fmaxf(m, logits[i]);
	m = block_max(m, red);
	float s = 0.f;
	for (int i = tid; i < vocab; i += kSampleThreads) s += expf((logits[i] - m) / temperature);
	const float sum = block_sum(s, red);
	// contiguous chunk per thread -> prefix over threads -> the first thread whose inclusive prefix reaches r walks its chunk
	const int chunk = (vocab + kSampleThreads - 1) / kSampleThreads;
	const int i0 = min(vocab, tid * chunk), i1 = min(vocab, i0 + chunk);
	float cs = 0.f;
	for (int i = i0; i < i1; ++i) cs += expf((logits[i] - m) / temperature) / sum;
	float incl = cs;
#pragma unroll
	for (int o = 1; o < 32; o <<= 1) {
		const float t = __shfl_up_sync(kFull, incl, o);
		if (lane >= o) incl += t;
	}
	if (lane == 31) wtot[wid] = incl;
	if (tid == 0) first = kSampleThreads;
	__syncthreads();
	float base = 0.f;
	for (int w = 0; w < wid; ++w) base += wtot[w];
	incl += base;
	const float r = r_in >= 0.f ? r_in : uniform_of(seed, ctl ? (unsigned long long)ctl[CTL_STEP] : counter);
	if (incl >= r && i1 > i0) atomicMin(&first, tid);
	__syncthreads();
	if (tid == (first < kSampleThreads ? first : 0)) {
		int tok = vocab - 1; // src/sampler.cpp:64
		if (first < kSampleThreads) {
			float cum = incl - cs;
			tok = i1 - 1; // rounding: the chunk's own running sum may stop a hair short of its prefix total
			for (int i = i0; i < i1; ++i) {
				cum += expf((logits[i] - m) / temperature) / sum;
				if (cum >= r) {
					tok = i;
					break;
				}
			}
		}
		if (token_out) *token_out = tok;
		if (ctl) {
			const int step = ctl[CTL_STEP];
			if (tokens_out) tokens_out[step] = tok;
			ctl[CTL_STEP] = step + 1;
			ctl[CTL_TOKEN] = tok;
			const int pos = ctl[CTL_POS] + 1;
			ctl[CTL_POS] = pos;
			derive_ring_ctl(ctl, pos, max_seq_len);
		}
	}
}

// probs[row] = softmax(logits[row, :])[targets[row]]; one CTA per row.
__global__ void __launch_bounds__(256) token_probs_kernel(const float *__restrict__ logits, int vocab, const int *__restrict__ targets,
                                                          float *__restrict__ probs) {
	__shared__ float red[33];
	const float *row = logits + (size_t)blockIdx.x * vocab;
	float m = -FLT_MAX;
	for (int i = threadIdx.x; i < vocab; i += blockDim.x) m = fmaxf(m, row[i]);
	m = block_max(m, red);
	float s = 0.f;
	for (int i = threadIdx.x; i < vocab; i += blockDim.x) s += expf(row[i] - m);
	s = block_sum(s, red);
	if (threadIdx.x == 0) {
		const int t = targets[blockIdx.x];
		probs[blockIdx.x] = (t >= 0 && t < vocab) ? expf(row[t] - m) / s : 0.f;
	}
}

} // namespace yalm

using namespace yalm;

extern "C" int yalm_sample_advance(const float *logits, int vocab, int *ctl, int max_seq_len, int *tokens_out,
                                   float temperature, unsigned long long seed, yalm_stream_t s) {
	if (!logits || !ctl || vocab <= 0 || max_seq_len <= KV_SINKS) return YALM_EINVAL;
	if (temperature == 0.f) return yalm_argmax_advance(logits, vocab, ctl, max_seq_len, tokens_out, s);
	if (!(temperature > 0.f)) return YALM_EINVAL;
	sample_kernel<<<1, kSampleThreads, 0, (cudaStream_t)s>>>(logits, vocab, temperature, -1.f, seed, ctl, max_seq_len, tokens_out, nullptr, 0ull);
	return (int)cudaGetLastError();
}

extern "C" int yalm_token_probs(const float *logits, int vocab, int n_rows, const int *targets, float *probs, yalm_stream_t s) {
	if (!logits || !targets || !probs || vocab <= 0 || n_rows < 1) return YALM_EINVAL;
	token_probs_kernel<<<n_rows, 256, 0, (cudaStream_t)s>>>(logits, vocab, targets, probs);
	return (int)cudaGetLastError();
}

// first-max arg-max of a device row into a device int (Sampler::sample_device at temperature 0)
__global__ void __launch_bounds__(1024) argmax_kernel(const float *__restrict__ logits, int vocab, int *token_out) {
	__shared__ float sv[32];
	__shared__ int si[32];
	float best = -FLT_MAX;
	int bi = 0x7fffffff;
	for (int i = threadIdx.x; i < vocab; i += blockDim.x) {
		const float v = logits[i];
		if (v > best) best = v, bi = i;
	}
	auto merge = [&]() {
		for (int o = 16; o > 0; o >>= 1) {
			const float ov = __shfl_xor_sync(kFull, best, o);
			const int oi = __shfl_xor_sync(kFull, bi, o);
			if (ov > best || (ov == best && oi < bi)) best = ov, bi = oi;
		}
	};
	merge();
	if ((threadIdx.x & 31) == 0) sv[threadIdx.x >> 5] = best, si[threadIdx.x >> 5] = bi;
	__syncthreads();
	if (threadIdx.x < 32) {
		best = sv[threadIdx.x], bi = si[threadIdx.x];
		merge();
		if (threadIdx.x == 0) *token_out = bi == 0x7fffffff ? 0 : bi;
	}
}

// One draw into *token_dev (device): r = uniform(seed, counter); temperature 0 = first-max arg-max.
extern "C" int yalm_sample_device1(const float *logits, int vocab, float temperature, unsigned long long seed,
                                   unsigned long long counter, int *token_dev, yalm_stream_t s) {
	if (!logits || !token_dev || vocab <= 0 || !(temperature >= 0.f)) return YALM_EINVAL;
	if (temperature == 0.f) argmax_kernel<<<1, 1024, 0, (cudaStream_t)s>>>(logits, vocab, token_dev);
	else sample_kernel<<<1, kSampleThreads, 0, (cudaStream_t)s>>>(logits, vocab, temperature, -1.f, seed, nullptr, 8, nullptr, token_dev, counter);
	return (int)cudaGetLastError();
}

// Host-pointer test entry: one draw with a caller-supplied r (Sampler::sample with std::rand() / RAND_MAX replaced by r).
extern "C" int yalm_sample_host(const float *logits, int vocab, float temperature, float r, int *token) {
	int n = 0;
	if (cudaGetDeviceCount(&n) != cudaSuccess || n == 0) return YALM_ENOGPU;
	if (!logits || !token || vocab <= 0 || !(temperature > 0.f) || r < 0.f) return YALM_EINVAL;
	float *d = nullptr;
	int *dt = nullptr;
	YALM_TRY(cudaMalloc(&d, sizeof(float) * vocab));
	YALM_TRY(cudaMalloc(&dt, sizeof(int)));
	YALM_TRY(cudaMemcpy(d, logits, sizeof(float) * vocab, cudaMemcpyHostToDevice));
	sample_kernel<<<1, kSampleThreads>>>(d, vocab, temperature, r, 0ull, nullptr, 8, nullptr, dt, 0ull);
	cudaError_t e = cudaMemcpy(token, dt, sizeof(int), cudaMemcpyDeviceToHost);
	cudaFree(d);
	cudaFree(dt);
	return (int)e;
}

extern "C" int yalm_token_probs_host(const float *logits, int vocab, int n_rows, const int *targets, float *probs) {
	int n = 0;
	if (cudaGetDeviceCount(&n) != cudaSuccess || n == 0) return YALM_ENOGPU;
	if (!logits || !targets || !probs || vocab <= 0 || n_rows < 1) return YALM_EINVAL;
	float *d = nullptr, *dp = nullptr;
	int *dt = nullptr;
	YALM_TRY(cudaMalloc(&d, sizeof(float) * (size_t)vocab * n_rows));
	YALM_TRY(cudaMalloc(&dp, sizeof(float) * n_rows));
	YALM_TRY(cudaMalloc(&dt, sizeof(int) * n_rows));
	YALM_TRY(cudaMemcpy(d, logits, sizeof(float) * (size_t)vocab * n_rows, cudaMemcpyHostToDevice));
	YALM_TRY(cudaMemcpy(dt, targets, sizeof(int) * n_rows, cudaMemcpyHostToDevice));
	int rc = yalm_token_probs(d, vocab, n_rows, dt, dp, nullptr);
	cudaError_t e = cudaMemcpy(probs, dp, sizeof(float) * n_rows, cudaMemcpyDeviceToHost);
	cudaFree(d);
	cudaFree(dp);
	cudaFree(dt);
	return rc ? rc : (int)e;
}
