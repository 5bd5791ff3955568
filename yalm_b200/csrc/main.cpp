// build/main — the reference's command line (src/main.cpp:17-42, 290-429) over the B200 forward pass:
//
//   main <checkpoint> [-d cuda|cpu] [-m completion|passkey|perplexity] [-T context] [-i prompt | -f file]
//                     [-n steps] [-t temperature] [-l passkey position]
//
// Same flags, the same unique-prefix matching of mode / device names, the same stats blocks.  What differs is
// underneath: `-d cuda` is the only device (there is no CPU execution path in this library: `-d cpu` is refused), the
// next token is sampled on the device (4 bytes per token come back instead of vocab * 4), and perplexity scores whole
// prompt chunks per launch when the model supports the batched prefill path (Model::forward_batch).
#include "sampler.h"
#include "tokenizer.h"
#include "yalm_api.h"

#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <iostream>
#include <sstream>
#include <string>
#include <vector>

namespace {

uint64_t now_ms() {
	return (uint64_t)std::chrono::duration_cast<std::chrono::milliseconds>(std::chrono::system_clock::now().time_since_epoch()).count();
}

[[noreturn]] void usage() {
	fputs("Usage:   main <checkpoint> [options]\n"
	      "Example: main model.yalm -i \"Q: What is the meaning of life?\"\n"
	      "Options:\n"
	      "  -h Display this help message\n"
	      "  -d [cpu,cuda] which device to use (default - cuda; this build has no CPU path and refuses cpu)\n"
	      "  -m [completion,passkey,perplexity] which mode to run in (default - completion)\n"
	      "  -T <int> sliding window context length (0 - max)\n"
	      "\n"
	      "Perplexity mode options:\n"
	      "  Choose one:\n"
	      "    -i <string> input prompt\n"
	      "    -t <float> temperature (default - 1.0)\n"
	      "    -f <filepath> input file with prompt\n"
	      "Completion mode options:\n"
	      "  -n <int>    number of steps to run for in completion mode, default 256. 0 = max_seq_len, -1 = infinite\n"
	      "  Choose one:\n"
	      "    -i <string> input prompt\n"
	      "    -f <filepath> input file with prompt\n"
	      "Passkey mode options:\n"
	      "  -n <int>    number of junk lines to insert (default - 250)\n"
	      "  -l <int>    passkey position (-1 - random)\n",
	      stderr);
	exit(1);
}

struct Options {
	std::string checkpoint, device = "cuda", mode = "completion", prompt, prompt_path;
	int context = 0, num_steps = 256, n_junk = 250, passkey_pos = -1;
	float temperature = 1.0f;
};

// `given` selects the one of `names` it is a prefix of, first match in list order (src/main.cpp:330-357)
bool pick(const std::string &given, std::initializer_list<const char *> names, std::string *out) {
	for (const char *n : names)
		if (std::string(n).compare(0, given.size(), given) == 0 && given.size() <= strlen(n)) {
			*out = n;
			return true;
		}
	return false;
}

Options parse(int argc, char **argv) {
	Options o;
	if (argc < 2) usage();
	o.checkpoint = argv[1];
	for (int i = 2; i < argc; i += 2) {
		if (i + 1 >= argc || argv[i][0] != '-' || strlen(argv[i]) != 2) usage();
		const char *v = argv[i + 1];
		switch (argv[i][1]) {
		case 'm':
			if (!pick(v, {"completion", "passkey", "perplexity"}, &o.mode)) usage();
			break;
		case 'd':
			if (!pick(v, {"cpu", "cuda"}, &o.device)) usage();
			break;
		case 'i': o.prompt = v; break;
		case 'f': o.prompt_path = v; break;
		case 't': o.temperature = std::stof(v); break;
		case 'T': o.context = std::stoi(v); break;
		case 'l': o.passkey_pos = std::stoi(v); break;
		case 'n': o.num_steps = o.n_junk = std::stoi(v); break;
		default: usage();
		}
	}
	return o;
}

// Everything the three modes share: checkpoint, model + state on the GPU, sampler, tokenizer, one warm-up forward.
struct Session {
	YALMData data;
	std::unique_ptr<Model> model;
	std::unique_ptr<InferenceState> state;
	std::unique_ptr<Sampler> sampler;
	std::unique_ptr<Tokenizer> tokenizer;

	Session(const Options &o) {
		if (data.from_file(o.checkpoint) != 0) {
			std::cerr << "Error: could not load checkpoint " << o.checkpoint << std::endl;
			exit(1);
		}
		model = std::make_unique<Model>(data, o.context);
		state = std::make_unique<InferenceState>(model->config);
		sampler = std::make_unique<Sampler>(model->config, now_ms());
		tokenizer = std::make_unique<Tokenizer>(data);
		std::cout << "Model active bytes with full context window: " << model->config->active_bytes(model->config->max_seq_len) << std::endl;
		std::cout << "Using CUDA" << std::endl;
		model->cuda();
		state->cuda();
		state->set_host_logits(false); // sampling happens on the device
		model->forward(*state, 0, 0);  // warm-up, as the reference does (src/main.cpp:66-70)
	}

	std::vector<int> encode(const std::string &prompt, bool print_ids) {
		const uint64_t t0 = now_ms();
		std::vector<int> enc = tokenizer->encode(prompt, true);
		const double s = (now_ms() - t0) / 1000.0;
		if (print_ids) std::cout << tokenizer->encoding_to_debug_string(enc) << std::endl;
		printf("Encoding stats: (%zu tokens, throughput: %.5gtok/s, latency: %.5gs/tok, total: %.5gs)\n\n", enc.size(), enc.size() / s,
		       s / enc.size(), s);
		return enc;
	}
};

void run_completion(const Options &o) {
	Session S(o);
	Model &model = *S.model;
	InferenceState &state = *S.state;
	int num_steps = o.num_steps == 0 ? model.config->max_seq_len : o.num_steps; // -n 0: the full context
	std::vector<int> enc = S.encode(o.prompt, true);
	const uint64_t start = now_ms();
	size_t read_bytes = 0;
	// hydrate the KV cache with the prompt; only the last position needs logits (src/main.cpp:91-97)
	for (size_t pos = 0; pos < enc.size(); ++pos) {
		model.forward(state, enc[pos], (int)pos, pos + 1 == enc.size() ? InferenceMode::OUTPUT_LOGITS : InferenceMode::HYDRATE_KV_CACHE);
		read_bytes += model.config->active_bytes(pos);
	}
	const uint64_t hydrated = now_ms();
	for (int i = 0; i < num_steps || num_steps == -1; ++i) { // src/main.cpp:102-112
		const int tok = S.sampler->sample_device(state, o.temperature);
		std::cout << S.tokenizer->decode_one(enc.back(), tok) << std::flush;
		enc.push_back(tok);
		if (tok == S.tokenizer->eos_id || tok == S.tokenizer->eot_id) break;
		model.forward(state, tok, (int)enc.size() - 1);
		read_bytes += model.config->active_bytes(enc.size() - 1);
	}
	std::cout << "\n" << std::endl;
	const double el = (now_ms() - start) / 1000.0;
	printf("Generation stats:\n  %zu tokens\n  throughput: %.5gtok/s\n  latency: %.5gs/tok\n  hydrate: %.5gs\n  bandwidth: %.5gGB/s\n  total: %.5gs\n\n",
	       enc.size(), enc.size() / el, el / enc.size(), (hydrated - start) / 1000.0, read_bytes / 1e9 / el, el);
}

void run_perplexity(const Options &o) {
	Session S(o);
	Model &model = *S.model;
	InferenceState &state = *S.state;
	std::vector<int> enc = S.encode(o.prompt, true);
	if (enc.size() < 2) {
		std::cerr << "Error: perplexity needs at least two tokens" << std::endl;
		exit(1);
	}
	const size_t N = enc.size() - 1;
	double sum_lp = 0.0, ss_lp = 0.0;
	size_t read_bytes = 0;
	const uint64_t start = now_ms();
	// log p(token[pos + 1] | token[0..pos]) for every position (src/main.cpp:174-184); the probability is taken on
	// the device, one float per position crosses PCIe
	std::vector<float> probs(N);
	size_t done = 0;
	if (model.supports_batch()) {
		// batched prefill: a chunk of positions per pass, dense GEMMs on the tensor cores
		const int chunk = model.batch_capacity();
		while (done < N) {
			const int n = (int)std::min<size_t>(chunk, N - done);
			std::cout << "\r Computing perplexity..." << done + n << "/" << N << std::flush;
			if (model.forward_batch(state, enc.data() + done, n, (int)done, enc.data() + done + 1, probs.data() + done) != 0) break;
			for (int i = 0; i < n; ++i) read_bytes += model.config->active_bytes(done + i) / n; // the weights are read once per chunk
			done += n;
		}
	}
	for (size_t pos = done; pos < N; ++pos) {
		std::cout << "\r Computing perplexity..." << pos + 1 << "/" << N << std::flush;
		model.forward(state, enc[pos], (int)pos);
		read_bytes += model.config->active_bytes(pos);
		probs[pos] = S.sampler->sample_prob_device(enc[pos + 1], state);
	}
	for (size_t pos = 0; pos < N; ++pos) {
		const double lp = std::log((double)probs[pos]);
		sum_lp += lp, ss_lp += lp * lp;
	}
	std::cout << std::endl;
	const double el = (now_ms() - start) / 1000.0;
	const double ppl = std::exp(-sum_lp / N), err = ppl * std::sqrt((ss_lp - sum_lp * sum_lp / N) / N / N);
	printf("Stats:\n  %zu tokens\n  perplexity: %.5g \xC2\xB1 %.5g\n  throughput: %.5gtok/s\n  latency: %.5gs/tok\n  bandwidth: %.5gGB/s\n  total: %.5gs\n\n", N,
	       ppl, err, N / el, el / N, read_bytes / 1e9 / el, el);
}

void run_passkey(const Options &o) {
	Session S(o);
	Model &model = *S.model;
	InferenceState &state = *S.state;
	const std::string prefix = "There is an important info hidden inside a lot of irrelevant text. "
	                           "Find it and memorize them. I will quiz you about the important information there.";
	const std::string suffix = " What is the pass key? The pass key is";
	const int passkey = std::rand() % 50000 + 1;
	const int at = o.passkey_pos == -1 ? std::rand() % o.n_junk : o.passkey_pos;
	std::string prompt = prefix;
	for (int i = 0; i < o.n_junk; ++i) {
		if (i == at) prompt += " The pass key is " + std::to_string(passkey) + ". Remember it. " + std::to_string(passkey) + " is the pass key.";
		prompt += " The grass is green. The sky is blue. The sun is yellow. Here we go. There and back again.";
	}
	prompt += suffix;
	std::vector<int> enc = S.encode(prompt, false);
	printf("Passkey test:\n  prompt: %zu tokens\n  passkey: %d\n  passkey token index: ~%d\n\n", enc.size(), passkey,
	       (int)((float)at / o.n_junk * enc.size()));
	const size_t N = enc.size();
	for (size_t pos = 0; pos < N; ++pos) {
		std::cout << "\r Running passkey test..." << pos + 1 << "/" << N << std::flush;
		model.forward(state, enc[pos], (int)pos, pos + 1 == N ? InferenceMode::OUTPUT_LOGITS : InferenceMode::HYDRATE_KV_CACHE);
	}
	std::cout << std::endl << suffix << std::flush;
	for (size_t pos = N; pos < N + 16; ++pos) { // at most 16 greedy steps (src/main.cpp:268-287)
		const int tok = S.sampler->sample_device(state, 0.f);
		std::cout << S.tokenizer->decode_one(enc.back(), tok) << std::flush;
		enc.push_back(tok);
		if (tok == S.tokenizer->eos_id || tok == S.tokenizer->eot_id) break;
		model.forward(state, tok, (int)pos);
	}
	std::cout << std::endl;
}

} // namespace

int main(int argc, char **argv) {
	Options o = parse(argc, argv);
	if (o.mode == "completion" || o.mode == "perplexity") {
		if ((o.prompt.empty() ? 0 : 1) + (o.prompt_path.empty() ? 0 : 1) != 1) usage();
		if (!o.prompt_path.empty()) {
			std::ifstream f(o.prompt_path);
			if (!f.is_open()) {
				std::cerr << "Error: could not open file " << o.prompt_path << std::endl;
				return 1;
			}
			std::stringstream buf;
			buf << f.rdbuf();
			o.prompt = buf.str();
		}
	} else if (o.passkey_pos != -1 && (o.passkey_pos >= o.n_junk || o.passkey_pos < 0)) {
		std::cerr << "Error: passkey position must be between 0 and " << o.n_junk - 1 << std::endl;
		return 1;
	}
	if (o.device != "cuda") {
		std::cerr << "Error: -d cpu is not available: yalm_b200 implements the CUDA (sm_100a) forward pass only and has no CPU fallback" << std::endl;
		return 1;
	}
	std::cout << "Using checkpoint: " << o.checkpoint << std::endl;
	if (o.mode == "completion") run_completion(o);
	else if (o.mode == "passkey") run_passkey(o);
	else run_perplexity(o);
	return 0;
}
