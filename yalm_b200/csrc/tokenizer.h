// Host-side text <-> token ids for the `main` CLI.  Same public surface as the reference's Tokenizer
// (src/tokenizer.h:68-72: encode / decode_one / encoding_to_debug_string, bos_id / eos_id / eot_id,
// byte_fallback_start) and the same behaviour — greedy longest-prefix match over `tokenizer.tokens`, byte fallback
// through the "<0x00>".."<0xFF>" pieces, leading-space strip after BOS (src/tokenizer.cpp:44-94) — on a different
// structure: the pieces are sorted once and a match is a binary search narrowed one byte at a time, instead of a trie of
// hash maps.  Not on the measured path (benchmarks feed token ids).
#pragma once

#include "codec.h"

#include <string>
#include <vector>

struct Tokenizer {
	std::vector<std::string> vocab; // index = token id
	int bos_id = -1, eos_id = -1, eot_id = -1;
	int byte_fallback_start = -1; // id of "<0x00>", -1 if the vocabulary has no byte pieces

	explicit Tokenizer(const YALMData &data);

	std::vector<int> encode(const std::string &text, bool encode_bos) const;
	std::string decode_one(int prev_token, int token) const;
	std::string encoding_to_debug_string(const std::vector<int> &encoding) const;

private:
	std::vector<int> _sorted; // token ids ordered by piece (bytewise); duplicates keep the highest id last
};
