// `.yalm` checkpoint container: u64 LE header length | JSON header | raw tensor bytes (the safetensors
// layout the reference parses in src/codec.cpp:116-175).  Same public types as the reference's
// src/codec.h (DType, Tensor, YALMData, dtype_size, dtype_to_string) so Model code reads the same;
// the JSON header is parsed by a small built-in parser instead of a vendored JSON library, and an
// image that is already in memory can be opened without touching the file system.
#pragma once

#include <array>
#include <cstddef>
#include <cstdint>
#include <map>
#include <string>
#include <unordered_map>

typedef uint16_t f16_t;

// Values match the reference enum (src/codec.h:15-25); they are also the YALM_* dtype codes of the C ABI.
enum class DType { F32 = 0, F16 = 1, BF16 = 2, F8E5M2 = 3, F8E4M3 = 4, I32 = 5, I16 = 6, I8 = 7, U8 = 8 };

std::string dtype_to_string(DType dtype);
size_t dtype_size(DType dtype);

struct Tensor {
	std::string name;
	DType dtype = DType::F32;
	std::array<int, 4> shape = {0, 0, 0, 0};
	void *data = nullptr;
	size_t size = 0; // bytes
};

// `__metadata__`: every value is a string (src/model.cpp:17-75 reads them with std::stoi/stof).
struct Metadata {
	std::map<std::string, std::string> kv;
	bool contains(const std::string &k) const { return kv.count(k) != 0; }
	const std::string &at(const std::string &k) const;                       // aborts with a message if absent
	std::string value(const std::string &k, const std::string &dflt) const { // like json::value
		auto it = kv.find(k);
		return it == kv.end() ? dflt : it->second;
	}
};

struct YALMData {
	void *data = nullptr;
	size_t size = 0;
	Metadata metadata;
	std::unordered_map<std::string, Tensor> tensors;

	YALMData() = default;
	YALMData(const YALMData &) = delete;
	YALMData &operator=(const YALMData &) = delete;
	~YALMData();

	// 0 on success, -1 on failure (reference convention, src/codec.cpp:116-175).
	int from_file(const std::string &filename);
	// Parse an image that lives in caller-owned memory (not copied, not freed).
	int from_memory(const void *image, size_t bytes);

	const Tensor *find(const std::string &name) const {
		auto it = tensors.find(name);
		return it == tensors.end() ? nullptr : &it->second;
	}

private:
	bool _mapped = false;
	int parse();
};
