// TEST INFRASTRUCTURE.  A reference-style C++ caller compiled against yalm_b200/csrc/yalm_api.h: the differential
// kernel tests of the reference's own test binary (src/test.cpp:148-206 — matmul 16x256 fp32, mha 16 heads / 8 KV heads
// / head_dim 16 / kv_len 4, ffn 256 <-> 256 GELU; inputs from std::default_random_engine seeds 0..3; abs tolerance 1e-4),
// with the CUDA side = this library's matmul_cuda / mha_cuda / ffn_cuda and the CPU side = the oracle (oracle/yalm_oracle.h)
// instead of the reference's *_cpu functions.  Also drives Model / InferenceState / Block through the C++ API the way
// src/main.cpp does.   usage: test_kernels [model.yalm]
#include "yalm_api.h"

#include "../../oracle/yalm_oracle.h"

#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <random>
#include <string>
#include <vector>

static void fill_random(float *data, size_t n, unsigned long seed, float scale = 1.0f) {
	std::default_random_engine gen(seed);
	std::normal_distribution<float> dist(0.0, 1.0);
	for (size_t i = 0; i < n; i++) data[i] = dist(gen) * scale;
}
static void fill_random(f16_t *data, size_t n, unsigned long seed, float scale = 1.0f) {
	std::default_random_engine gen(seed);
	std::normal_distribution<float> dist(0.0, 1.0);
	for (size_t i = 0; i < n; i++) data[i] = yo_float_to_half(dist(gen) * scale);
}
static int check(const std::vector<float> &got, const std::vector<float> &want, const char *what, float eps = 1e-4f) {
	for (size_t i = 0; i < got.size(); i++)
		if (!(std::fabs(got[i] - want[i]) <= eps)) {
			fprintf(stderr, "FAIL %s: [%zu] got %g want %g\n", what, i, got[i], want[i]);
			return 1;
		}
	printf("ok %s\n", what);
	return 0;
}

int main(int argc, char **argv) {
	const int head_dim = 16, n_heads = 16, dim = head_dim * n_heads, hidden_dim = dim, n_kv_heads = 8, max_seq_len = 4, kv_len = 4;
	int bad = 0;
	{ // matmul
		std::vector<float> w(dim * head_dim), x(dim), cpu(head_dim), gpu(head_dim);
		fill_random(w.data(), w.size(), 0);
		fill_random(x.data(), x.size(), 1);
		yo_matmul(cpu.data(), x.data(), w.data(), nullptr, dim, head_dim, YO_F32);
		matmul_cuda(gpu.data(), x.data(), w.data(), dim, head_dim);
		bad += check(gpu, cpu, "matmul");
	}
	{ // mha
		std::vector<f16_t> kb(max_seq_len * n_kv_heads * head_dim), vb(kb.size());
		fill_random(kb.data(), kb.size(), 0);
		fill_random(vb.data(), vb.size(), 1);
		std::vector<float> q(n_heads * head_dim), att_c(n_heads * max_seq_len), att_g(att_c.size()), out_c(q.size()), out_g(q.size());
		fill_random(q.data(), q.size(), 2);
		yo_mha(out_c.data(), att_c.data(), kb.data(), vb.data(), q.data(), head_dim, kv_len, max_seq_len, n_heads, n_kv_heads);
		mha_cuda(out_g.data(), att_g.data(), kb.data(), vb.data(), q.data(), head_dim, kv_len, max_seq_len, n_heads, n_kv_heads);
		bad += check(att_g, att_c, "mha att");
		bad += check(out_g, out_c, "mha xout");
	}
	{ // ffn
		std::vector<float> x(dim), w1(dim * hidden_dim), w2(w1.size()), w3(w1.size()), cpu(dim), gpu(dim);
		fill_random(x.data(), x.size(), 0);
		fill_random(w1.data(), w1.size(), 1, 1.0f / sqrtf(dim));
		fill_random(w2.data(), w2.size(), 2, 1.0f / sqrtf(hidden_dim));
		fill_random(w3.data(), w3.size(), 3, 1.0f / sqrtf(dim));
		yo_ffn(cpu.data(), x.data(), w1.data(), w2.data(), w3.data(), nullptr, nullptr, nullptr, hidden_dim, dim, YO_GELU, YO_F32);
		ffn_cuda(gpu.data(), x.data(), w1.data(), w2.data(), w3.data(), hidden_dim, dim, ActivationType::GELU);
		bad += check(gpu, cpu, "ffn");
	}
	if (argc > 1) { // the run_completion call sequence (src/main.cpp:44-112) through the C++ API, greedy
		YALMData data;
		if (data.from_file(argv[1]) != 0) return 2;
		Model model(data, 0);
		InferenceState state(model.config);
		model.cuda();
		state.cuda();
		model.forward(state, 0, 0);
		const int prompt[] = {1, 17, 300, 42, 5};
		for (int pos = 0; pos < 5; pos++)
			model.forward(state, prompt[pos], pos, pos + 1 == 5 ? InferenceMode::OUTPUT_LOGITS : InferenceMode::HYDRATE_KV_CACHE);
		printf("tokens:");
		for (int i = 0; i < 8; i++) {
			const int tok = yo_argmax(state.logits(), model.config->vocab_size);
			printf(" %d", tok);
			model.forward(state, tok, 5 + i);
		}
		printf("\n");
		// Block-level accessors compile and point at device memory
		const Block &b = *model.blocks[0];
		printf("block0 wq %s kc %s\n", b.wq<f16_t>() ? "set" : "null", b.key_cache() ? "set" : "null");
	}
	return bad ? 1 : 0;
}
