// decode_mega_kernel instantiations for YALM_F32 weights, default variant (one translation unit per weight type and kernel
// variant: build time, and the default kernel carries none of the optional paths).
#include "decode_mega.cuh"

namespace yalm {
namespace mega {
const void *kernel_f32(int head_dim, int g, bool tp) {
	set_fault_pointers(fault_record(), fault_timeout_cycles());
	return kernel_for<YALM_F32, 0>(head_dim, g, tp);
}
} // namespace mega
} // namespace yalm
