#include "loader.h"

#include <cuda_runtime.h>

#include <algorithm>
#include <cstdlib>
#include <cstring>

namespace yalm {

DeviceArena::~DeviceArena() {
	if (_base) cudaFree(_base);
}

int DeviceArena::reserve(size_t bytes) {
	if (_base) return (int)cudaErrorInvalidValue;
	_cap = padded(bytes);
	_used = 0;
	return (int)cudaMalloc((void **)&_base, std::max<size_t>(_cap, 256));
}

void *DeviceArena::take(size_t bytes) {
	const size_t n = padded(bytes);
	if (!_base || _used + n > _cap) return nullptr;
	void *p = _base + _used;
	_used += n;
	return p;
}

Uploader::Uploader(size_t staging_bytes) {
	if (!staging_bytes) {
		const char *env = getenv("YALM_LOAD_STAGING_MB");
		staging_bytes = (size_t)(env && atoi(env) > 0 ? atoi(env) : 256) << 20;
	}
	_cap = staging_bytes;
	cudaStreamCreateWithFlags(&_stream, cudaStreamNonBlocking);
	for (int i = 0; i < 2; ++i) {
		if (cudaHostAlloc((void **)&_buf[i], _cap, cudaHostAllocDefault) != cudaSuccess) {
			// no pinned memory to be had: put() degrades to synchronous pageable copies (still one arena)
			cudaGetLastError();
			_buf[i] = nullptr;
			if (i == 1) cudaFreeHost(_buf[0]), _buf[0] = nullptr;
			break;
		}
		cudaEventCreateWithFlags(&_done[i], cudaEventDisableTiming);
	}
}

Uploader::~Uploader() {
	finish();
	for (int i = 0; i < 2; ++i) {
		if (_buf[i]) cudaFreeHost(_buf[i]);
		if (_done[i]) cudaEventDestroy(_done[i]);
	}
	if (_stream) cudaStreamDestroy(_stream);
}

// rows x width bytes (row r at src + r * spitch) -> staging buffer (host threads) -> dst (copy engine)
int Uploader::stage_and_send(char *dst, const char *src, size_t spitch, size_t width, size_t rows) {
	const int b = _next;
	_next ^= 1;
	if (_busy[b]) {
		cudaError_t e = cudaEventSynchronize(_done[b]);
		if (e != cudaSuccess) return (int)e;
		_busy[b] = false;
	}
	char *stage = _buf[b];
	const size_t total = width * rows;
	if (width == spitch || rows == 1) {
		// one contiguous range: split it over the host threads (page-in of the mmap happens here, in parallel)
		const size_t piece = 4u << 20;
		const long n = (long)((total + piece - 1) / piece);
#pragma omp parallel for schedule(dynamic, 1)
		for (long i = 0; i < n; ++i) {
			const size_t o = (size_t)i * piece;
			memcpy(stage + o, src + o, std::min(piece, total - o));
		}
	} else {
#pragma omp parallel for schedule(static)
		for (long r = 0; r < (long)rows; ++r) memcpy(stage + (size_t)r * width, src + (size_t)r * spitch, width);
	}
	cudaError_t e = cudaMemcpyAsync(dst, stage, total, cudaMemcpyHostToDevice, _stream);
	if (e != cudaSuccess) return (int)e;
	e = cudaEventRecord(_done[b], _stream);
	_busy[b] = true;
	_moved += total;
	return (int)e;
}

int Uploader::put(void *dst, const void *host, size_t spitch, size_t offset, size_t width, size_t height) {
	if (!dst || !host || !width || !height) return 0;
	const char *src = static_cast<const char *>(host) + offset;
	char *d = static_cast<char *>(dst);
	if (!_buf[0]) { // pageable fallback
		_moved += width * height;
		if (width == spitch || height == 1) return (int)cudaMemcpy(d, src, width * height, cudaMemcpyHostToDevice);
		return (int)cudaMemcpy2D(d, width, src, spitch, width, height, cudaMemcpyHostToDevice);
	}
	if (width == spitch || height == 1) { // contiguous: staging-buffer-sized bites
		const size_t total = width * height;
		for (size_t o = 0; o < total; o += _cap) {
			int rc = stage_and_send(d + o, src + o, _cap, std::min(_cap, total - o), 1);
			if (rc) return rc;
		}
		return 0;
	}
	if (width > _cap) { // a single piece larger than a buffer (not a shape the model produces): piece by piece
		for (size_t r = 0; r < height; ++r) {
			int rc = put(d + r * width, src + r * spitch, width, 0, width, 1);
			if (rc) return rc;
		}
		return 0;
	}
	const size_t rows_per = std::max<size_t>(1, _cap / width);
	for (size_t r = 0; r < height; r += rows_per) {
		int rc = stage_and_send(d + r * width, src + r * spitch, spitch, width, std::min(rows_per, height - r));
		if (rc) return rc;
	}
	return 0;
}

int Uploader::zero(void *dst, size_t bytes) {
	return (int)cudaMemsetAsync(dst, 0, bytes, _stream);
}

int Uploader::finish() {
	cudaError_t e = _stream ? cudaStreamSynchronize(_stream) : cudaSuccess;
	_busy[0] = _busy[1] = false;
	return (int)e;
}

} // namespace yalm
