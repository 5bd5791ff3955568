#include "codec.h"

#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>

#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

std::string dtype_to_string(DType dtype) {
	static const char *names[] = {"F32", "F16", "BF16", "F8_E5M2", "F8_E4M3", "I32", "I16", "I8", "U8"};
	int i = static_cast<int>(dtype);
	return (i >= 0 && i < 9) ? names[i] : "UNKNOWN";
}

size_t dtype_size(DType dtype) {
	static const size_t sizes[] = {4, 2, 2, 1, 1, 4, 2, 1, 1};
	int i = static_cast<int>(dtype);
	return (i >= 0 && i < 9) ? sizes[i] : 0;
}

static bool dtype_from_string(const std::string &s, DType *out) {
	for (int i = 0; i < 9; ++i)
		if (dtype_to_string(static_cast<DType>(i)) == s) {
			*out = static_cast<DType>(i);
			return true;
		}
	return false;
}

const std::string &Metadata::at(const std::string &k) const {
	auto it = kv.find(k);
	if (it == kv.end()) {
		fprintf(stderr, "FATAL: missing metadata key: %s\n", k.c_str());
		abort();
	}
	return it->second;
}

// ---- a JSON reader for exactly the subset a safetensors header uses ---------------------------------
// { "name": { "dtype": "F16", "shape": [a,b], "data_offsets": [s,e] }, "__metadata__": { "k": "v" } }
namespace {
struct Cursor {
	const char *p, *end;
	bool ok = true;
	void ws() {
		while (p < end && (*p == ' ' || *p == '\n' || *p == '\t' || *p == '\r')) ++p;
	}
	bool eat(char c) {
		ws();
		if (p < end && *p == c) {
			++p;
			return true;
		}
		return false;
	}
	bool expect(char c) {
		if (!eat(c)) ok = false;
		return ok;
	}
	std::string str() {
		std::string out;
		ws();
		if (p >= end || *p != '"') {
			ok = false;
			return out;
		}
		++p;
		while (p < end && *p != '"') {
			if (*p == '\\' && p + 1 < end) {
				++p;
				switch (*p) {
				case 'n': out += '\n'; break;
				case 't': out += '\t'; break;
				case 'r': out += '\r'; break;
				case 'b': out += '\b'; break;
				case 'f': out += '\f'; break;
				case 'u': { // \uXXXX -> UTF-8 (BMP only; enough for metadata strings)
					if (p + 4 >= end) { ok = false; return out; }
					unsigned cp = (unsigned)strtoul(std::string(p + 1, p + 5).c_str(), nullptr, 16);
					p += 4;
					if (cp < 0x80) out += (char)cp;
					else if (cp < 0x800) out += (char)(0xC0 | (cp >> 6)), out += (char)(0x80 | (cp & 0x3F));
					else out += (char)(0xE0 | (cp >> 12)), out += (char)(0x80 | ((cp >> 6) & 0x3F)), out += (char)(0x80 | (cp & 0x3F));
					break;
				}
				default: out += *p;
				}
				++p;
			} else {
				out += *p++;
			}
		}
		if (p >= end) ok = false;
		else ++p;
		return out;
	}
	bool number(double *v) {
		ws();
		char *e = nullptr;
		*v = strtod(p, &e);
		if (e == p) return ok = false;
		p = e;
		return true;
	}
	bool int_array(std::vector<long long> *out) {
		if (!expect('[')) return false;
		if (eat(']')) return true;
		do {
			double v;
			if (!number(&v)) return false;
			if (v != (double)(long long)v) return ok = false;
			out->push_back((long long)v);
		} while (eat(','));
		return expect(']');
	}
	// skip any value (used for keys we do not know)
	void skip() {
		ws();
		if (p >= end) { ok = false; return; }
		if (*p == '"') { str(); return; }
		if (*p == '{' || *p == '[') {
			char open = *p, close = open == '{' ? '}' : ']';
			++p;
			if (eat(close)) return;
			do {
				if (open == '{') { str(); expect(':'); }
				skip();
			} while (ok && eat(','));
			expect(close);
			return;
		}
		while (p < end && *p != ',' && *p != '}' && *p != ']') ++p; // number / true / false / null
	}
};
} // namespace

int YALMData::parse() {
	if (size < sizeof(uint64_t)) return -1;
	uint64_t json_size;
	memcpy(&json_size, data, sizeof(json_size));
	if (json_size == 0 || json_size > size - sizeof(uint64_t)) return -1;
	const char *json_ptr = static_cast<const char *>(data) + sizeof(uint64_t);
	char *bytes_ptr = static_cast<char *>(data) + sizeof(uint64_t) + json_size;
	const size_t bytes_size = size - sizeof(uint64_t) - json_size;

	Cursor c{json_ptr, json_ptr + json_size};
	if (!c.expect('{')) return -1;
	if (!c.eat('}')) {
		do {
			std::string key = c.str();
			if (!c.expect(':')) return -1;
			if (key == "__metadata__") {
				if (!c.expect('{')) return -1;
				if (!c.eat('}')) {
					do {
						std::string k = c.str();
						if (!c.expect(':')) return -1;
						c.ws();
						if (c.p < c.end && *c.p == '"') metadata.kv[k] = c.str();
						else c.skip();
					} while (c.ok && c.eat(','));
					if (!c.expect('}')) return -1;
				}
				continue;
			}
			Tensor t;
			t.name = key;
			std::string dtype_str;
			std::vector<long long> shape, offs;
			if (!c.expect('{')) return -1;
			do {
				std::string f = c.str();
				if (!c.expect(':')) return -1;
				if (f == "dtype") dtype_str = c.str();
				else if (f == "shape") c.int_array(&shape);
				else if (f == "data_offsets") c.int_array(&offs);
				else c.skip();
			} while (c.ok && c.eat(','));
			if (!c.expect('}')) return -1;
			if (!dtype_from_string(dtype_str, &t.dtype)) {
				fprintf(stderr, "bad dtype\n");
				return -1;
			}
			if (shape.size() > 4) fprintf(stderr, "shape exceeds 4 dimensions\n");
			size_t numel = 1;
			for (size_t i = 0; i < shape.size() && i < 4; ++i) {
				if (shape[i] < 0 || shape[i] > 0x7fffffffLL) {
					fprintf(stderr, "bad shape\n");
					return -1;
				}
				t.shape[i] = (int)shape[i];
				numel *= (size_t)shape[i];
			}
			if (offs.size() != 2 || offs[0] < 0 || offs[1] <= offs[0] || (size_t)offs[1] > bytes_size) {
				fprintf(stderr, "bad offsets\n");
				return -1;
			}
			t.data = bytes_ptr + offs[0];
			t.size = (size_t)(offs[1] - offs[0]);
			if (numel * dtype_size(t.dtype) != t.size) {
				fprintf(stderr, "bad size\n");
				return -1;
			}
			tensors[key] = t;
		} while (c.ok && c.eat(','));
		if (!c.expect('}')) return -1;
	}
	return c.ok ? 0 : -1;
}

int YALMData::from_file(const std::string &filename) {
	fprintf(stderr, "loading data from file: %s\n", filename.c_str());
	int fd = open(filename.c_str(), O_RDONLY);
	if (fd == -1) return -1;
	struct stat st;
	if (fstat(fd, &st) != 0) {
		close(fd);
		return -1;
	}
	size = st.st_size;
	data = mmap(nullptr, size, PROT_READ | PROT_WRITE, MAP_PRIVATE, fd, 0);
	if (data == MAP_FAILED) {
		data = nullptr;
		close(fd);
		return -1;
	}
	posix_fadvise(fd, 0, size, POSIX_FADV_SEQUENTIAL);
	close(fd);
	_mapped = true;
	if (parse() != 0) {
		munmap(data, size);
		data = nullptr, _mapped = false;
		tensors.clear();
		return -1;
	}
	return 0;
}

int YALMData::from_memory(const void *image, size_t bytes) {
	data = const_cast<void *>(image);
	size = bytes;
	_mapped = false;
	if (parse() != 0) {
		data = nullptr;
		tensors.clear();
		return -1;
	}
	return 0;
}

YALMData::~YALMData() {
	if (_mapped && data) munmap(data, size);
}
