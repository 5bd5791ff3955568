// ./build/test — kernel benchmarks with the reference's command line (src/test.cpp:361-383):
//   test -k <matmul|matmul-wide|mha|ffn|stream>   (also accepted: -b = matmul, -b2 = matmul-wide, -bk <name>)
// Each benchmark runs the reference's shapes (Mistral-7B: dim 4096, hidden 14336, 32/8 heads x 128, kv_len 4096)
// on the device through the C ABI and prints mean microseconds per launch and the achieved GB/s.
// Correctness of the same kernels is covered by tests/ (pytest -m gpu), not here.
#include <cstdio>
#include <cstring>
#include <string>

#include "yalm_b200.h"

static int usage(const char *argv0) {
	std::fprintf(stderr, "usage: %s -k <matmul|matmul-wide|mha|ffn|stream> | -b | -b2 | -bk <name>\n", argv0);
	return 1;
}

int main(int argc, char **argv) {
	std::string name;
	for (int i = 1; i < argc; ++i) {
		const std::string a = argv[i];
		if (a == "-b") name = "matmul";
		else if (a == "-b2") name = "matmul-wide";
		else if ((a == "-k" || a == "-bk") && i + 1 < argc) name = argv[++i];
		else return usage(argv[0]);
	}
	if (name.empty()) return usage(argv[0]);
	double usec = 0.0, bytes = 0.0;
	const int rc = yalm_bench_op(name.c_str(), /*iters=*/20, /*warmup=*/3, &usec, &bytes);
	if (rc != 0) {
		std::fprintf(stderr, "%s: %s (%d)\n", name.c_str(), yalm_error_string(rc), rc);
		return 1;
	}
	std::printf("%s: %.2f us/launch, %.1f GB/s (%.1f MB per launch)\n", name.c_str(), usec, bytes / usec / 1e3, bytes / 1e6);
	return 0;
}
