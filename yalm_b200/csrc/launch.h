// Host-side launch sizing: persistent grids = (#SMs x resident CTAs per SM), capped by the task count.
#pragma once

#include <cuda_runtime.h>

#include <algorithm>
#include <map>
#include <mutex>
#include <utility>

namespace yalm {

struct DeviceInfo {
	int device = -1;
	int sms = 0;
	int max_smem_optin = 0;
};

inline const DeviceInfo &device_info() {
	static thread_local DeviceInfo info;
	int dev = 0;
	cudaGetDevice(&dev);
	if (info.device != dev) {
		info.device = dev;
		cudaDeviceGetAttribute(&info.sms, cudaDevAttrMultiProcessorCount, dev);
		cudaDeviceGetAttribute(&info.max_smem_optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev);
	}
	return info;
}

// Resident CTAs per SM for (kernel, block, smem); raises the kernel's dynamic-smem limit on demand.
// Results are cached: the occupancy query is not free and this sits on the per-token path when the
// ops are used outside a CUDA graph.
inline int resident_ctas(const void *kernel, int block, size_t smem, cudaError_t *err) {
	static std::mutex mu;
	static std::map<std::pair<const void *, std::pair<int, size_t>>, int> cache;
	static std::map<const void *, size_t> smem_limit;
	std::lock_guard<std::mutex> lock(mu);
	*err = cudaSuccess;
	if (smem > 48 * 1024) {
		size_t &lim = smem_limit[kernel];
		if (smem > lim) {
			*err = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
			if (*err != cudaSuccess) return 0;
			lim = smem;
		}
	}
	auto key = std::make_pair(kernel, std::make_pair(block, smem));
	auto it = cache.find(key);
	if (it != cache.end()) return it->second;
	int occ = 0;
	*err = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kernel, block, smem);
	if (*err != cudaSuccess) return 0;
	cache[key] = occ;
	return occ;
}

// Grid for `tasks` warp-tasks with `warps_per_cta` warps per CTA.
inline int persistent_grid(const void *kernel, int block, size_t smem, long tasks, cudaError_t *err) {
	int occ = resident_ctas(kernel, block, smem, err);
	if (*err != cudaSuccess) return 0;
	if (occ < 1) {
		*err = cudaErrorLaunchOutOfResources;
		return 0;
	}
	const int warps_per_cta = block / 32;
	long need = (tasks + warps_per_cta - 1) / warps_per_cta;
	long cap = (long)device_info().sms * occ;
	return (int)std::max(1L, std::min(need, cap));
}

} // namespace yalm
