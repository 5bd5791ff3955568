// Tensor-parallel plumbing: one process per GPU, NCCL over NVLink 5 / NVSwitch.
// NCCL is bound at run time (dlopen of the libnccl.so.2 that torch bundles, or the system one) so that
// the single-GPU build has no link-time dependency on it.  Only the five calls the decode path needs.
#include "tp.h"

#include <dlfcn.h>

#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>

namespace yalm {
namespace {

typedef struct ncclComm *ncclComm_t;
struct ncclUniqueId_ {
	char internal[128];
};
constexpr int kNcclFloat32 = 7, kNcclSum = 0; // nccl.h: ncclDataType_t / ncclRedOp_t

struct Api {
	void *lib = nullptr;
	int (*GetUniqueId)(ncclUniqueId_ *) = nullptr;
	int (*CommInitRank)(ncclComm_t *, int, ncclUniqueId_, int) = nullptr;
	int (*AllReduce)(const void *, void *, size_t, int, int, ncclComm_t, cudaStream_t) = nullptr;
	int (*AllGather)(const void *, void *, size_t, int, ncclComm_t, cudaStream_t) = nullptr;
	int (*CommDestroy)(ncclComm_t) = nullptr;
	const char *(*GetErrorString)(int) = nullptr;
};

Api g_api;
ncclComm_t g_comm = nullptr;
int g_rank = 0, g_world = 1;

bool load_api() {
	if (g_api.lib) return true;
	const char *env = getenv("YALM_NCCL_LIB");
	const char *names[] = {env, "libnccl.so.2", "libnccl.so"};
	for (const char *n : names) {
		if (!n) continue;
		g_api.lib = dlopen(n, RTLD_NOW | RTLD_GLOBAL);
		if (g_api.lib) break;
	}
	if (!g_api.lib) {
		fprintf(stderr, "yalm_b200: cannot load NCCL (%s); set YALM_NCCL_LIB\n", dlerror());
		return false;
	}
#define YALM_SYM(field, name)                                                                                 \
	*(void **)(&g_api.field) = dlsym(g_api.lib, name);                                                        \
	if (!g_api.field) {                                                                                       \
		fprintf(stderr, "yalm_b200: NCCL symbol %s missing\n", name);                                         \
		return false;                                                                                         \
	}
	YALM_SYM(GetUniqueId, "ncclGetUniqueId")
	YALM_SYM(CommInitRank, "ncclCommInitRank")
	YALM_SYM(AllReduce, "ncclAllReduce")
	YALM_SYM(AllGather, "ncclAllGather")
	YALM_SYM(CommDestroy, "ncclCommDestroy")
	YALM_SYM(GetErrorString, "ncclGetErrorString")
#undef YALM_SYM
	return true;
}

int check(int rc, const char *what) {
	if (rc != 0) {
		fprintf(stderr, "yalm_b200: %s failed: %s\n", what, g_api.GetErrorString ? g_api.GetErrorString(rc) : "?");
		return -100 - rc;
	}
	return 0;
}

} // namespace

int tp_rank() { return g_rank; }
int tp_world() { return g_world; }
bool tp_active() { return g_comm != nullptr && g_world > 1; }

int tp_allreduce_sum(const float *send, float *recv, size_t count, cudaStream_t st) {
	if (!tp_active()) return YALM_EINVAL;
	return check(g_api.AllReduce(send, recv, count, kNcclFloat32, kNcclSum, g_comm, st), "ncclAllReduce");
}

int tp_allgather(const float *send, float *recv, size_t count_per_rank, cudaStream_t st) {
	if (!tp_active()) return YALM_EINVAL;
	return check(g_api.AllGather(send, recv, count_per_rank, kNcclFloat32, g_comm, st), "ncclAllGather");
}

} // namespace yalm

using namespace yalm;

extern "C" int yalm_tp_unique_id(void *id128) {
	if (!id128 || !load_api()) return YALM_EINVAL;
	ncclUniqueId_ id;
	int rc = check(g_api.GetUniqueId(&id), "ncclGetUniqueId");
	if (rc) return rc;
	memcpy(id128, &id, sizeof(id));
	return 0;
}

extern "C" int yalm_tp_init(const void *id128, int rank, int world) {
	if (!id128 || rank < 0 || world < 1 || rank >= world) return YALM_EINVAL;
	if (world == 1) {
		g_rank = 0, g_world = 1;
		return 0;
	}
	if (!load_api()) return YALM_EINVAL;
	ncclUniqueId_ id;
	memcpy(&id, id128, sizeof(id));
	int rc = check(g_api.CommInitRank(&g_comm, world, id, rank), "ncclCommInitRank");
	if (rc) return rc;
	g_rank = rank, g_world = world;
	return 0;
}

extern "C" void yalm_tp_finalize(void) {
	if (g_comm) g_api.CommDestroy(g_comm);
	g_comm = nullptr;
	g_rank = 0, g_world = 1;
}

// ---- shard arithmetic (pure; shared by the loader and the CPU tests) ---------------------------------------
namespace yalm {

bool tp_shards(const TpDims &d, int size) {
	return size >= 1 && d.n_kv_heads % size == 0 && d.n_heads % size == 0 && d.hidden % size == 0 && d.vocab % size == 0;
}

TpSlice tp_slice(const TpDims &d, int tensor, int rank, int size) {
	const size_t B = d.bytes_per_weight, T = size, r = rank, dim = d.dim, hid = d.hidden;
	const size_t q_dim = (size_t)d.n_heads * d.head_dim, kv_dim = (size_t)d.n_kv_heads * d.head_dim;
	const size_t lq = q_dim / T, lkv = kv_dim / T, lhid = hid / T, lv = (size_t)d.vocab / T;
	switch (tensor) {
	// column-parallel: whole output rows -> one contiguous piece (w1/w3: per expert, spitch = one expert)
	case kTpWq: return {lq * dim * B, r * lq * dim * B, lq * dim * B, 1};
	case kTpWk:
	case kTpWv: return {lkv * dim * B, r * lkv * dim * B, lkv * dim * B, 1};
	case kTpW1:
	case kTpW3: return {hid * dim * B, r * lhid * dim * B, lhid * dim * B, 1};
	case kTpWcls: return {lv * dim * B, r * lv * dim * B, lv * dim * B, 1};
	// row-parallel: a column range of every output row
	case kTpWo: return {q_dim * B, r * lq * B, lq * B, dim};
	case kTpW2: return {hid * B, r * lhid * B, lhid * B, dim};
	default: return {0, 0, 0, 0};
	}
}

} // namespace yalm

// dims[7]: dim, hidden, head_dim, n_heads, n_kv_heads, vocab, bytes per weight.  out[4]: spitch, offset, width, height.
extern "C" int yalm_tp_shard_plan(const int *dims, int tensor, int rank, int size, long long *out) {
	if (!dims || !out || tensor < 0 || tensor >= yalm::kTpTensors || rank < 0 || rank >= size) return YALM_EINVAL;
	const yalm::TpDims d{dims[0], dims[1], dims[2], dims[3], dims[4], dims[5], dims[6]};
	if (!yalm::tp_shards(d, size)) return YALM_EINVAL;
	const yalm::TpSlice s = yalm::tp_slice(d, tensor, rank, size);
	out[0] = (long long)s.spitch, out[1] = (long long)s.offset, out[2] = (long long)s.width, out[3] = (long long)s.height;
	return 0;
}
