#include "tokenizer.h"

#include <algorithm>

Tokenizer::Tokenizer(const YALMData &data) {
	bos_id = std::stoi(data.metadata.at("bos_token_id"));
	eos_id = std::stoi(data.metadata.at("eos_token_id"));
	const Tensor *t = data.find("tokenizer.tokens");
	if (t) { // NUL-separated UTF-8 pieces, id = position in the list (src/tokenizer.cpp:11-19)
		const char *p = static_cast<const char *>(t->data), *end = p + t->size;
		while (p < end) {
			const char *q = p;
			while (q < end && *q != '\0') ++q;
			vocab.emplace_back(p, q - p);
			p = q + 1;
		}
	}
	for (size_t i = 0; i < vocab.size(); ++i) {
		if (vocab[i] == "<0x00>") byte_fallback_start = (int)i;
		else if (vocab[i] == "<|eot_id|>" || vocab[i] == "<|end|>" || vocab[i] == "<|im_end|>") eot_id = (int)i;
	}
	_sorted.resize(vocab.size());
	for (size_t i = 0; i < vocab.size(); ++i) _sorted[i] = (int)i;
	// bytewise order (std::string compares as unsigned char); equal pieces: ascending id, so the LAST of a run of
	// duplicates is the id the reference's trie ends up holding (later insertions overwrite, src/tokenizer.cpp:32-41)
	std::stable_sort(_sorted.begin(), _sorted.end(), [&](int a, int b) { return vocab[a] < vocab[b]; });
}

std::string Tokenizer::decode_one(int prev_token, int token) const {
	const std::string &piece = vocab[token];
	if (prev_token == bos_id && !piece.empty() && piece[0] == ' ') return piece.substr(1); // sentencepiece strips it after BOS
	if (byte_fallback_start >= 0 && token >= byte_fallback_start && token - byte_fallback_start < 256)
		return std::string(1, (char)(token - byte_fallback_start));
	return piece;
}

std::vector<int> Tokenizer::encode(const std::string &text, bool encode_bos) const {
	std::vector<int> out;
	if (encode_bos) out.push_back(bos_id);
	const size_t n = _sorted.size();
	for (size_t i = 0; i < text.size();) {
		// [lo, hi) = pieces that start with text[i, i + l); grows l one byte at a time
		size_t lo = 0, hi = n, l = 0, best_l = 0;
		int best = -1;
		while (i + l < text.size() && lo < hi) {
			const unsigned char c = (unsigned char)text[i + l];
			auto key = [&](int id) -> int { // byte l of the piece, -1 if the piece ends there
				const std::string &s = vocab[id];
				return l < s.size() ? (int)(unsigned char)s[l] : -1;
			};
			lo = std::partition_point(_sorted.begin() + lo, _sorted.begin() + hi, [&](int id) { return key(id) < (int)c; }) - _sorted.begin();
			hi = std::partition_point(_sorted.begin() + lo, _sorted.begin() + hi, [&](int id) { return key(id) <= (int)c; }) - _sorted.begin();
			if (lo >= hi) break;
			++l;
			// pieces of exactly l bytes sort first in the range; the last duplicate wins
			size_t e = lo;
			while (e < hi && vocab[_sorted[e]].size() == l) ++e;
			if (e > lo) best = _sorted[e - 1], best_l = l;
		}
		if (best < 0) {
			if (byte_fallback_start >= 0) out.push_back((int)(unsigned char)text[i] + byte_fallback_start);
			i += 1;
		} else {
			out.push_back(best);
			i += best_l;
		}
	}
	return out;
}

std::string Tokenizer::encoding_to_debug_string(const std::vector<int> &encoding) const {
	std::string s;
	for (int id : encoding) {
		if (id == bos_id) s += "[<s>:" + std::to_string(id) + "]";
		else if (id == eos_id) s += "[</s>:" + std::to_string(id) + "]";
		else s += "[" + vocab[id] + ":" + std::to_string(id) + "]";
	}
	return s;
}
