// decode_mega_kernel instantiations for YALM_F16 weights, variant 2: tensor-core GEMV (decode_mega.cuh "tensor-core GEMV").
#include "decode_mega.cuh"

namespace yalm {
namespace mega {
const void *kernel_f16_mma(int head_dim, int g, bool tp) {
	set_fault_pointers(fault_record(), fault_timeout_cycles());
	return kernel_for<YALM_F16, 2>(head_dim, g, tp);
}
} // namespace mega
} // namespace yalm
