// Tensor-parallel plumbing: one process per GPU, NCCL over NVLink 5 / NVSwitch.
// NCCL is bound at run time (dlopen of the libnccl.so.2 that torch bundles, or the system one) so that
// the single-GPU build has no link-time dependency on it.  Only the five calls the decode path needs.
#include "tp.h"
#include "mega.h"

#include <dlfcn.h>

#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

namespace yalm {
namespace {

typedef struct ncclComm *ncclComm_t;
struct ncclUniqueId_ {
	char internal[128];
};
constexpr int kNcclFloat32 = 7, kNcclInt32 = 2, kNcclChar = 0, kNcclSum = 0, kNcclMin = 3; // nccl.h: ncclDataType_t / ncclRedOp_t

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


// ---- one-shot all-reduce over NVLink peer memory ---------------------------------------------------------------
// The decode path all-reduces one fp32[dim] vector (16-32 KB) twice per layer: latency, not bandwidth.  Every rank
// owns a small region that all peers map (CUDA IPC); a single-CTA kernel pushes the rank's partial vector into slot
// [rank] of every peer's region with plain stores over NVLink, raises a flag there, waits for the world's flags in
// its own region and sums the slots in rank order — so every rank computes the same bits, run after run.  Slots and
// flags are double-buffered by the parity of a per-rank call counter that lives on the device, which keeps the
// kernel replayable inside a CUDA graph (no host-side sequence number in the launch arguments).
constexpr int kPeerMaxWorld = 8;
constexpr int kPeerCapFloats = 8192;                    // largest vector (dim) the one-shot path takes
constexpr size_t kPeerFlagsOff = 256, kPeerSlotsOff = 1024;
constexpr size_t kPeerOneShotBytes = kPeerSlotsOff + 2ull * kPeerMaxWorld * kPeerCapFloats * sizeof(float);
constexpr size_t kPeerMegaOff = (kPeerOneShotBytes + 1023) / 1024 * 1024; // the decode megakernel's area (mega.h: kTpOff*)
constexpr size_t kPeerRegionBytes = kPeerMegaOff + mega::kTpRegionBytes;
static_assert(kPeerMaxWorld == mega::kTpMaxWorld, "one world-size limit");
constexpr long long kPeerTimeoutCycles = 4000000000ll; // ~2 s: a lost peer traps instead of hanging the GPU

struct PeerPtrs {
	char *p[kPeerMaxWorld];
};

__global__ void __launch_bounds__(1024, 1) peer_allreduce_kernel(PeerPtrs pp, int world, int rank, const float *send,
                                                                  float *recv, int count) {
	char *mine = pp.p[rank];
	unsigned *seqp = reinterpret_cast<unsigned *>(mine);
	const unsigned seq = *reinterpret_cast<volatile unsigned *>(seqp) + 1; // calls so far + 1, same on every rank
	const unsigned par = seq & 1u;
	const int n4 = count / 4, cap4 = kPeerCapFloats / 4;
	const int tid = threadIdx.x;
	// 1. push this rank's partial into slot [par][rank] of every region
	for (int i = tid; i < n4; i += blockDim.x) {
		const float4 v = reinterpret_cast<const float4 *>(send)[i];
		for (int r = 0; r < world; ++r)
			reinterpret_cast<float4 *>(pp.p[r] + kPeerSlotsOff)[(size_t)(par * world + rank) * cap4 + i] = v;
	}
	__threadfence_system();
	__syncthreads();
	if (tid < world) {
		// 2. tell peer `tid` that slot [par][rank] is complete ...
		unsigned *flag = reinterpret_cast<unsigned *>(pp.p[tid] + kPeerFlagsOff) + par * kPeerMaxWorld + rank;
		asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(flag), "r"(seq) : "memory");
		// 3. ... and wait until rank `tid` has filled its slot here
		const unsigned *in = reinterpret_cast<const unsigned *>(mine + kPeerFlagsOff) + par * kPeerMaxWorld + tid;
		unsigned v;
		const long long t0 = clock64();
		do {
			asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(in) : "memory");
			if (clock64() - t0 > kPeerTimeoutCycles) __trap();
		} while (v != seq);
	}
	__syncthreads();
	// 4. sum the slots in rank order
	for (int i = tid; i < n4; i += blockDim.x) {
		float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
		for (int r = 0; r < world; ++r) {
			const float4 v = __ldcg(reinterpret_cast<const float4 *>(mine + kPeerSlotsOff) + (size_t)(par * world + r) * cap4 + i);
			acc.x += v.x, acc.y += v.y, acc.z += v.z, acc.w += v.w;
		}
		reinterpret_cast<float4 *>(recv)[i] = acc;
	}
	__syncthreads();
	if (tid == 0) *seqp = seq;
}

namespace {
struct Peer {
	bool ok = false;
	char *local = nullptr;
	PeerPtrs ptrs{};
} g_peer;

void peer_close() {
	for (int r = 0; r < kPeerMaxWorld; ++r)
		if (g_peer.ptrs.p[r] && g_peer.ptrs.p[r] != g_peer.local) cudaIpcCloseMemHandle(g_peer.ptrs.p[r]);
	if (g_peer.local) cudaFree(g_peer.local);
	g_peer = Peer{};
	cudaGetLastError();
}

// Map every rank's region.  Any failure on any rank -> all ranks fall back to NCCL (agreed with an all-reduce(min)).
void peer_open(int rank, int world) {
	const char *env = getenv("YALM_TP_PEER");
	int good = (!env || atoi(env) != 0) && world <= kPeerMaxWorld ? 1 : 0;
	cudaIpcMemHandle_t mine_h{};
	char *dev = nullptr;
	if (cudaMalloc(&dev, world * sizeof(cudaIpcMemHandle_t) + 16) != cudaSuccess) good = 0;
	if (good && cudaMalloc(&g_peer.local, kPeerRegionBytes) != cudaSuccess) good = 0;
	if (good && cudaMemset(g_peer.local, 0, kPeerRegionBytes) != cudaSuccess) good = 0;
	if (good && cudaIpcGetMemHandle(&mine_h, g_peer.local) != cudaSuccess) good = 0;
	cudaDeviceSynchronize();
	cudaGetLastError();
	std::vector<cudaIpcMemHandle_t> all(world);
	if (dev) {
		cudaMemcpy(dev + rank * sizeof(cudaIpcMemHandle_t), &mine_h, sizeof(mine_h), cudaMemcpyHostToDevice);
		if (g_api.AllGather(dev + rank * sizeof(cudaIpcMemHandle_t), dev, sizeof(cudaIpcMemHandle_t), kNcclChar, g_comm, 0) != 0) good = 0;
		cudaStreamSynchronize(0);
		cudaMemcpy(all.data(), dev, world * sizeof(cudaIpcMemHandle_t), cudaMemcpyDeviceToHost);
	}
	for (int r = 0; good && r < world; ++r) {
		if (r == rank) {
			g_peer.ptrs.p[r] = g_peer.local;
		} else {
			void *ptr = nullptr;
			if (cudaIpcOpenMemHandle(&ptr, all[r], cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) good = 0;
			g_peer.ptrs.p[r] = static_cast<char *>(ptr);
		}
	}
	cudaGetLastError();
	// agree: the one-shot path is used only if every rank mapped every region
	int *flag = dev ? reinterpret_cast<int *>(dev + world * sizeof(cudaIpcMemHandle_t)) : nullptr;
	int agreed = 0;
	if (flag) {
		cudaMemcpy(flag, &good, sizeof(int), cudaMemcpyHostToDevice);
		if (g_api.AllReduce(flag, flag, 1, kNcclInt32, kNcclMin, g_comm, 0) == 0) {
			cudaStreamSynchronize(0);
			cudaMemcpy(&agreed, flag, sizeof(int), cudaMemcpyDeviceToHost);
		}
	}
	if (dev) cudaFree(dev);
	if (agreed) g_peer.ok = true;
	else peer_close();
	if (rank == 0) fprintf(stderr, "yalm_b200: tensor-parallel all-reduce over %s\n", g_peer.ok ? "NVLink peer memory (one-shot kernel)" : "NCCL");
}
} // namespace

TpPeerView tp_peer_view() {
	TpPeerView v{};
	v.ok = g_peer.ok, v.world = g_world, v.rank = g_rank;
	for (int r = 0; r < kPeerMaxWorld; ++r) v.base[r] = g_peer.ok && g_peer.ptrs.p[r] ? g_peer.ptrs.p[r] + kPeerMegaOff : nullptr;
	return v;
}

int tp_rank() { return g_rank; }
int tp_world() { return g_world; }
bool tp_active() { return g_comm != nullptr && g_world > 1; }

int tp_allreduce_sum(const float *send, float *recv, size_t count, cudaStream_t st) {
	if (!tp_active()) return YALM_EINVAL;
	if (g_peer.ok && count <= (size_t)kPeerCapFloats && count % 4 == 0) {
		peer_allreduce_kernel<<<1, 1024, 0, st>>>(g_peer.ptrs, g_world, g_rank, send, recv, (int)count);
		return (int)cudaGetLastError();
	}
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
	peer_open(rank, world);
	return 0;
}

extern "C" void yalm_tp_finalize(void) {
	peer_close();
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
