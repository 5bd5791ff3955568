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

// Peer-mapped regions for the in-kernel all-reduce of the whole-token decode kernel (mega.h: kTpOff*): base[r] is
// rank r's region as mapped into this process, base[rank] the local one.  ok == false: no peer mapping (NCCL only).
struct TpPeerView {
	bool ok;
	int world, rank;
	char *base[8];
};
TpPeerView tp_peer_view();

// Which bytes of an unsharded (rows, cols) row-major tensor rank `rank` of `size` holds: `height` pieces of `width`
// bytes, piece i starting at offset + i * spitch.  Pure arithmetic (no device): shared by the loader and the tests.
// Column-parallel tensors (wq, wk, wv, w1, w3, wcls: split over output rows) are one contiguous piece; row-parallel
// ones (wo, w2: split over the input dimension) are one piece per output row.
enum TpTensor { kTpWq = 0, kTpWk, kTpWv, kTpWo, kTpW1, kTpW2, kTpW3, kTpWcls, kTpTensors };
struct TpSlice {
	size_t spitch, offset, width, height;
};
struct TpDims {
	int dim, hidden, head_dim, n_heads, n_kv_heads, vocab, bytes_per_weight;
};
bool tp_shards(const TpDims &d, int size); // does the model split `size` ways?
TpSlice tp_slice(const TpDims &d, int tensor, int rank, int size);
} // namespace yalm
