#include "sampler.h"

#include <cuda_runtime.h>

#include <cfloat>
#include <cmath>
#include <cstdio>
#include <cstdlib>

extern "C" int yalm_sample_device1(const float *logits, int vocab, float temperature, unsigned long long seed,
                                   unsigned long long counter, int *token_dev, yalm_stream_t s);

Sampler::Sampler(const std::shared_ptr<Config> config, uint64_t seed) : vocab_size(config->vocab_size), _seed(seed) {
	std::srand((unsigned)seed);
}

Sampler::~Sampler() {
	if (_scratch) cudaFree(_scratch);
}

static float row_max(const float *logits, int n) {
	float m = -FLT_MAX;
	for (int i = 0; i < n; ++i)
		if (logits[i] > m) m = logits[i];
	return m;
}

float Sampler::sample_prob(int index, const InferenceState &s) {
	const float *logits = s.logits();
	const float m = row_max(logits, vocab_size);
	float sum = 0;
	for (int i = 0; i < vocab_size; ++i) sum += expf(logits[i] - m);
	return expf(logits[index] - m) / sum;
}

int Sampler::sample_argmax(const InferenceState &s) {
	const float *logits = s.logits();
	int best = 0;
	float m = -FLT_MAX;
	for (int i = 0; i < vocab_size; ++i)
		if (logits[i] > m) m = logits[i], best = i;
	return best;
}

int Sampler::sample(const InferenceState &s, float temperature) {
	if (temperature == 0.0f) return sample_argmax(s);
	const float *logits = s.logits();
	const float m = row_max(logits, vocab_size);
	float sum = 0;
	for (int i = 0; i < vocab_size; ++i) sum += expf((logits[i] - m) / temperature);
	const float r = std::rand() / (float)RAND_MAX;
	float cum = 0;
	for (int i = 0; i < vocab_size; ++i) {
		cum += expf((logits[i] - m) / temperature) / sum;
		if (cum >= r) return i;
	}
	return vocab_size - 1;
}

static void die(int rc, const char *what) {
	if (rc != 0) {
		fprintf(stderr, "yalm_b200: %s failed: %s (%d)\n", what, yalm_error_string(rc), rc);
		abort();
	}
}

float Sampler::sample_prob_device(int index, const InferenceState &s) {
	if (s.device() != Device::CUDA) return sample_prob(index, s);
	if (!_scratch) die((int)cudaMalloc((void **)&_scratch, 16), "cudaMalloc");
	cudaStream_t st = s.stream();
	float p = 0.f;
	die((int)cudaMemcpyAsync(_scratch + 1, &index, sizeof(int), cudaMemcpyHostToDevice, st), "H2D");
	die(yalm_token_probs(s.logits_dev(), vocab_size, 1, _scratch + 1, reinterpret_cast<float *>(_scratch + 2), st), "yalm_token_probs");
	die((int)cudaMemcpyAsync(&p, _scratch + 2, sizeof(float), cudaMemcpyDeviceToHost, st), "D2H");
	die((int)cudaStreamSynchronize(st), "sync");
	return p;
}

int Sampler::sample_device(const InferenceState &s, float temperature) {
	if (s.device() != Device::CUDA) return sample(s, temperature);
	if (!_scratch) die((int)cudaMalloc((void **)&_scratch, 16), "cudaMalloc");
	cudaStream_t st = s.stream();
	int tok = 0;
	die(yalm_sample_device1(s.logits_dev(), vocab_size, temperature, _seed, _draws++, _scratch, st), "yalm_sample_device1");
	die((int)cudaMemcpyAsync(&tok, _scratch, sizeof(int), cudaMemcpyDeviceToHost, st), "D2H");
	die((int)cudaStreamSynchronize(st), "sync");
	return tok;
}
