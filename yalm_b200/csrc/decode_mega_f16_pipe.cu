// decode_mega_kernel instantiations for YALM_F16 weights, variant 1: pipelined hb (decode_mega.cuh "pipelined hb").
#include "decode_mega.cuh"

namespace yalm {
namespace mega {
const void *kernel_f16_pipe(int head_dim, int g, bool tp) {
	set_fault_pointers(fault_record(), fault_timeout_cycles());
	return kernel_for<YALM_F16, 1>(head_dim, g, tp);
}
} // namespace mega
} // namespace yalm
