// Checkpoint -> HBM upload path.  The reference does one cudaMalloc and one synchronous copy out of
// pageable mmap memory per tensor (src/model.cpp:185-211, 380-394; src/infer.cu:59-61): ~355 allocations for a 32-layer
// model and a PCIe stream that stalls on every page fault.  Here a model (or tensor-parallel shard) is ONE device
// allocation, and the bytes travel through two pinned staging buffers: host threads gather the next piece (the
// tensor-parallel row / column slice is cut right here, on the host side of the copy) out of the mmap into one buffer
// while the copy engine drains the other.
#pragma once

#include <cuda_runtime_api.h>

#include <cstddef>

namespace yalm {

// Bump allocator over one cudaMalloc; everything is 256-byte aligned (TMA bulk copies need 16).
class DeviceArena {
public:
	DeviceArena() = default;
	~DeviceArena();
	DeviceArena(const DeviceArena &) = delete;
	static size_t padded(size_t bytes) { return (bytes + 255) / 256 * 256; }
	int reserve(size_t bytes); // the one allocation; returns a cudaError_t
	void *take(size_t bytes);  // null when the reservation is exhausted (a sizing bug)
	size_t capacity() const { return _cap; }
	size_t used() const { return _used; }
	bool owns(const void *p) const { return p >= _base && p < _base + _cap; }

private:
	char *_base = nullptr;
	size_t _cap = 0, _used = 0;
};

// Pinned double-buffered host -> device stream.  put() returns as soon as the bytes are in a staging buffer and their
// DMA is queued; finish() waits for everything.  Not thread-safe (one loader thread per model).
class Uploader {
public:
	explicit Uploader(size_t staging_bytes = 0); // 0: YALM_LOAD_STAGING_MB or 256 MB per buffer
	~Uploader();
	Uploader(const Uploader &) = delete;
	// `height` pieces of `width` bytes, piece i read at host + offset + i * spitch  ->  dense at dst (device)
	int put(void *dst, const void *host, size_t spitch, size_t offset, size_t width, size_t height);
	int zero(void *dst, size_t bytes);
	int finish();
	size_t bytes_moved() const { return _moved; }
	bool pinned() const { return _buf[0] != nullptr; }

private:
	int stage_and_send(char *dst, const char *src, size_t spitch, size_t width, size_t rows);
	char *_buf[2] = {nullptr, nullptr};
	cudaEvent_t _done[2] = {nullptr, nullptr};
	bool _busy[2] = {false, false};
	cudaStream_t _stream = nullptr;
	size_t _cap = 0, _moved = 0;
	int _next = 0;
};

} // namespace yalm
