// The reference's Sampler (src/sampler.h:7-20) over the logits of an InferenceState — host methods with the reference's
// exact scan order (first maximum wins; running sum in index order), plus device twins that work on logits_dev() and
// bring one int / one float back instead of the whole vocabulary row.
#pragma once

#include "yalm_api.h"

#include <cstdint>
#include <memory>

struct Sampler {
	int vocab_size;

	Sampler(const std::shared_ptr<Config> config, uint64_t seed);
	~Sampler();
	Sampler(const Sampler &) = delete;

	// softmax(logits)[index] (src/sampler.cpp:11-25)
	float sample_prob(int index, const InferenceState &s);
	// index of the first maximum (src/sampler.cpp:27-38)
	int sample_argmax(const InferenceState &s);
	// draw with temperature; 0 = arg-max (src/sampler.cpp:40-65)
	int sample(const InferenceState &s, float temperature = 1.0);

	// Same three on the device logits of `s` (valid after an OUTPUT_LOGITS forward, also with set_host_logits(false)).
	float sample_prob_device(int index, const InferenceState &s);
	int sample_device(const InferenceState &s, float temperature = 1.0);

private:
	uint64_t _seed, _draws = 0;
	int *_scratch = nullptr; // device: {token, target}, {prob}
};
