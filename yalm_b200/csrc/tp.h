// Tensor-parallel collectives used by the decode path (see tp.cpp).
#pragma once

#include <cuda_runtime_api.h>
#include <stddef.h>

#include "../../include/yalm_b200.h"

namespace yalm {
int tp_rank();
int tp_world();
bool tp_active();
// recv = sum over ranks of send (fp32), fixed NCCL ring/tree order -> run-to-run deterministic.
int tp_allreduce_sum(const float *send, float *recv, size_t count, cudaStream_t st);
int tp_allgather(const float *send, float *recv, size_t count_per_rank, cudaStream_t st);
} // namespace yalm
