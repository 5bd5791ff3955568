// The seven memory / stream helpers the reference exports from src/infer.cu:59-90 (declared
// src/model.h:33-39): same names, same meaning, same error convention (stderr + abort()).
#include <cuda_runtime.h>

#include <cstdio>
#include <cstdlib>

#include "../../include/yalm_b200.h"

#define YALM_ABORT_ON(x)                                                                                      \
	do {                                                                                                      \
		cudaError_t err__ = (x);                                                                              \
		if (err__ != cudaSuccess) {                                                                           \
			fprintf(stderr, "CUDA error in %s at %s:%d: %s (%s=%d)\n", __FUNCTION__, __FILE__, __LINE__,       \
			        cudaGetErrorString(err__), cudaGetErrorName(err__), (int)err__);                          \
			abort();                                                                                          \
		}                                                                                                     \
	} while (0)

extern "C" void *upload_cuda(void *host, size_t size) {
	void *device = nullptr;
	YALM_ABORT_ON(cudaMalloc(&device, size));
	YALM_ABORT_ON(cudaMemcpy(device, host, size, cudaMemcpyHostToDevice));
	return device;
}

extern "C" void *download_cuda_raw(void *device, size_t size) {
	void *host = nullptr;
	YALM_ABORT_ON(cudaMallocHost(&host, size));
	YALM_ABORT_ON(cudaMemcpy(host, device, size, cudaMemcpyDeviceToHost));
	return host;
}

extern "C" void register_cuda_host(void *host, size_t size) {
	YALM_ABORT_ON(cudaHostRegister(host, size, cudaHostRegisterDefault));
}

extern "C" void free_cuda(void *device) {
	YALM_ABORT_ON(cudaFree(device));
}

extern "C" void unregister_cuda_host(void *host) {
	YALM_ABORT_ON(cudaHostUnregister(host));
}

extern "C" void set_cuda_device(int device) {
	YALM_ABORT_ON(cudaSetDevice(device));
}

extern "C" void init_cuda_stream(yalm_stream_t *stream) {
	YALM_ABORT_ON(cudaStreamCreateWithFlags(reinterpret_cast<cudaStream_t *>(stream), cudaStreamNonBlocking));
}

extern "C" const char *yalm_error_string(int code) {
	switch (code) {
	case 0: return "ok";
	case YALM_EINVAL: return "invalid argument or unsupported shape/dtype";
	case YALM_ENOGPU: return "no CUDA device (yalm_b200 has no CPU fallback)";
	case YALM_EFORMAT: return "bad .yalm container";
	}
	if (code > 0) return cudaGetErrorString(static_cast<cudaError_t>(code));
	return "communication error";
}

extern "C" const char *yalm_version(void) {
	return "yalm_b200 0.1 (sm_100a)";
}
