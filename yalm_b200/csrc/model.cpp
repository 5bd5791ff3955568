// Host side of the forward pass: configuration, scratch state, weight binding / sharding / upload,
// op sequencing and CUDA-graph capture.  See yalm_api.h for how this maps onto the reference API.
#include "yalm_api.h"

#include "loader.h"
#include "mega.h"
#include "tp.h"

#include <cuda_runtime.h>

#include <algorithm>
#include <cassert>
#include <cfloat>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <functional>
#include <iostream>

extern "C" int yalm_attn_generic(float *out, float *att, const uint16_t *key_cache, const uint16_t *value_cache,
                                 const float *q, int n_heads, int n_kv_heads, int head_dim, int max_seq_len,
                                 int kv_len, const int *ctl, yalm_stream_t s);
extern "C" int yalm_attn_is_fast_shape(int n_heads, int n_kv_heads, int head_dim);

#define CUDA_CHECK(x)                                                                                         \
	do {                                                                                                      \
		cudaError_t err__ = (x);                                                                              \
		if (err__ != cudaSuccess) {                                                                           \
			fprintf(stderr, "CUDA error in %s at %s:%d: %s (%s=%d)\n", __FUNCTION__, __FILE__, __LINE__,       \
			        cudaGetErrorString(err__), cudaGetErrorName(err__), (int)err__);                          \
			if (yalm::mega::describe_fault()[0]) fprintf(stderr, "yalm_b200: %s\n", yalm::mega::describe_fault()); \
			abort();                                                                                          \
		}                                                                                                     \
	} while (0)

// op-level failures on the hot path follow the reference convention too: report and abort
#define OP_CHECK(x)                                                                                           \
	do {                                                                                                      \
		int rc__ = (x);                                                                                       \
		if (rc__ != 0) {                                                                                      \
			fprintf(stderr, "yalm_b200: %s failed at %s:%d: %s (%d)\n", #x, __FILE__, __LINE__,               \
			        yalm_error_string(rc__), rc__);                                                           \
			abort();                                                                                          \
		}                                                                                                     \
	} while (0)

static inline size_t align_up(size_t v, size_t a) {
	return (v + a - 1) / a * a;
}

// ------------------------------------------------------------------------------------------------
// Config
// ------------------------------------------------------------------------------------------------
void Config::from_yalm(YALMData &yalm, int context) {
	const Metadata &md = yalm.metadata;
	auto geti = [&](const char *k) { return std::stoi(md.at(k)); };
	dim = geti("dim");
	hidden_dim = geti("hidden_dim");
	head_dim = geti("head_dim");
	n_layers = geti("n_layers");
	n_heads = geti("n_heads");
	n_kv_heads = geti("n_kv_heads");
	vocab_size = geti("vocab_size");
	n_experts = md.contains("n_experts") ? geti("n_experts") : 0;
	n_experts_active = md.contains("n_experts_active") ? geti("n_experts_active") : 0;
	// Reference rule (src/model.cpp:31-36): clamp to 4096 unless a context is given, which replaces it.
	max_seq_len = std::min(geti("max_seq_len"), 4096);
	if (context) max_seq_len = context;
	rope_theta = std::stof(md.at("rope_theta"));
	rotary_dim = geti("rotary_dim");
	norm_eps = std::stof(md.value("norm_eps", "1e-5"));
	std::string act_str = md.value("act_type", "gelu");
	if (act_str == "silu") act = ActivationType::SILU;
	else {
		if (act_str != "gelu") std::cerr << "unsupported act_type, defaulting to gelu" << std::endl;
		act = ActivationType::GELU;
	}
	if (md.value("norm_type", "rmsnorm") != "rmsnorm")
		std::cerr << "unsupported norm_type, defaulting to rmsnorm" << std::endl;
	norm_type = LayerNormType::RMSNorm;
	qkv_clip = md.contains("qkv_clip") ? std::stof(md.at("qkv_clip")) : FLT_MAX;
	std::string dtype = md.at("dtype");
	if (dtype == "fp32") weight_dtype = DType::F32;
	else if (dtype == "fp16") weight_dtype = DType::F16;
	else if (dtype == "fp8") {
		// The reference maps "fp8" to F8E5M2 and then reinterprets the bytes as fp16;
		// there is no arithmetic to reproduce.  We define fp8 = e4m3 + per-row scales and say which
		// flavour the file holds from the tensor dtype itself.
		const Tensor *t = yalm.find("model.layers.0.attn.wq.weight");
		weight_dtype = (t && t->dtype == DType::F8E5M2) ? DType::F8E5M2 : DType::F8E4M3;
	} else {
		std::cerr << "FATAL: unsupported dtype: " << dtype << std::endl;
		assert(false);
		abort();
	}
}

size_t Config::active_bytes(size_t pos) const {
	const size_t ws = dtype_size(weight_dtype);
	const size_t q_dim = (size_t)n_heads * head_dim, kv_dim = (size_t)n_kv_heads * head_dim;
	size_t per_block = 2 * (size_t)dim * sizeof(float);
	per_block += (2 * q_dim + 2 * kv_dim) * dim * ws;
	if (n_experts > 0) per_block += ((size_t)n_experts * dim + (size_t)n_experts_active * 3 * dim * hidden_dim) * ws;
	else per_block += (size_t)3 * dim * hidden_dim * ws;
	const size_t kv_len = std::min<size_t>(max_seq_len, pos + 1);
	per_block += 2 * kv_len * kv_dim * sizeof(f16_t);
	return (size_t)dim * ws + (size_t)n_layers * per_block + (size_t)dim * sizeof(float) + (size_t)vocab_size * dim * ws;
}

// ------------------------------------------------------------------------------------------------
// CudaGraph
// ------------------------------------------------------------------------------------------------
void CudaGraph::launch(cudaStream_t s) {
	CUDA_CHECK(cudaGraphLaunch(instance, s));
}
void CudaGraph::destroy() {
	if (instance) cudaGraphExecDestroy(instance);
	if (graph) cudaGraphDestroy(graph);
	instance = nullptr, graph = nullptr, is_created = false;
}

// ------------------------------------------------------------------------------------------------
// InferenceState
// ------------------------------------------------------------------------------------------------
InferenceState::InferenceState(const std::shared_ptr<Config> config) : _config(config) {
	assert(config);
	const Config &c = *config;
	// Host scratch, as the reference allocates it (src/model.cpp:267-284); replaced by device memory in cuda().
	_x = new float[c.dim]();
	_xb = new float[c.dim]();
	_xb2 = new float[std::max(c.dim, c.n_heads * c.head_dim)]();
	const int k_act = std::max(1, c.n_experts_active);
	_hb = new float[(size_t)k_act * c.hidden_dim]();
	_hb2 = new float[c.hidden_dim]();
	_q = new float[c.n_heads * c.head_dim]();
	_k = new float[c.n_kv_heads * c.head_dim]();
	_v = new float[c.n_kv_heads * c.head_dim]();
	_att = new float[(size_t)c.n_heads * c.max_seq_len]();
	_logits = new float[c.vocab_size]();
	// Unlike the reference (src/model.cpp:279-283, the dense-model null store of SURVEY.md P4) the
	// expert buffers always exist.
	_moe_weights = new float[std::max(1, c.n_experts)]();
	_active_experts = new int[k_act]();
	_active_experts_weights = new float[k_act]();
}

InferenceState::~InferenceState() {
	_hydrate_graph.destroy();
	_output_graph.destroy();
	_greedy_graph.destroy();
	_device_graph.destroy();
	_output_dev_graph.destroy();
	_profile_graph.destroy();
	for (cudaEvent_t e : _ffn_events) cudaEventDestroy(e);
	if (_device == Device::CPU) {
		delete[] _x;
		delete[] _xb;
		delete[] _xb2;
		delete[] _hb;
		delete[] _hb2;
		delete[] _q;
		delete[] _k;
		delete[] _v;
		delete[] _att;
		delete[] _moe_weights;
		delete[] _active_experts;
		delete[] _active_experts_weights;
		delete[] _logits;
	} else {
		cudaFree(_arena);
		cudaFreeHost(_logits);
		if (_stream) cudaStreamDestroy(_stream);
	}
}

void InferenceState::enable_ffn_timing(bool on) {
	if (on && _ffn_events.empty()) {
		_ffn_events.resize(2 * (size_t)_config->n_layers);
		for (auto &e : _ffn_events) CUDA_CHECK(cudaEventCreate(&e));
	}
	_ffn_timing = on;
}

void InferenceState::cuda() {
	if (_device != Device::CPU) return;
	const Config &c = *_config;
	delete[] _x;
	delete[] _xb;
	delete[] _xb2;
	delete[] _hb;
	delete[] _hb2;
	delete[] _q;
	delete[] _k;
	delete[] _v;
	delete[] _att;
	delete[] _moe_weights;
	delete[] _active_experts;
	delete[] _active_experts_weights;
	delete[] _logits;
	_device = Device::CUDA;
	init_cuda_stream(reinterpret_cast<yalm_stream_t *>(&_stream));

	const int lh = c.local_heads(), lkv = c.local_kv_heads(), lhid = c.local_hidden();
	const int k_act = std::max(1, c.n_experts_active);
	const bool generic_attn = !yalm_attn_is_fast_shape(lh, lkv, c.head_dim);
	_tokens_cap = 1 << 16;
	int sms = 0, dev = 0;
	CUDA_CHECK(cudaGetDevice(&dev));
	CUDA_CHECK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
	struct Slot {
		void **ptr;
		size_t bytes;
	};
	Slot slots[] = {
	    {(void **)&_x, sizeof(float) * c.dim},
	    {(void **)&_xb, sizeof(float) * c.dim}, // TP: partial sums before the all-reduce
	    {(void **)&_xb2, sizeof(float) * std::max(c.dim, lh * c.head_dim)},
	    {(void **)&_hb, sizeof(float) * (size_t)k_act * lhid},
	    {(void **)&_hb2, sizeof(float) * (size_t)lhid},
	    {(void **)&_q, sizeof(float) * lh * c.head_dim},
	    {(void **)&_k, sizeof(float) * lkv * c.head_dim},
	    {(void **)&_v, sizeof(float) * lkv * c.head_dim},
	    {(void **)&_att, generic_attn ? sizeof(float) * (size_t)lh * c.max_seq_len : 16},
	    {(void **)&_moe_weights, sizeof(float) * std::max(1, c.n_experts)},
	    {(void **)&_active_experts, sizeof(int) * k_act},
	    {(void **)&_active_experts_weights, sizeof(float) * k_act},
	    {(void **)&_logits_dev, sizeof(float) * c.vocab_size},
	    {(void **)&_ctl, sizeof(int) * YALM_CTL_INTS},
	    {(void **)&_attn_ws, std::max<size_t>(256, yalm_attn_workspace_bytes(lh, lkv, c.head_dim))},
	    {(void **)&_tokens_dev, sizeof(int) * (size_t)_tokens_cap},
	    {(void **)&_mega_acc, sizeof(float) * (size_t)sms * lh * c.head_dim},
	    {(void **)&_mega_ml, sizeof(float) * (size_t)sms * lh * 2},
	    {(void **)&_mega_counters, sizeof(unsigned) * (size_t)std::max(1, lkv)},
	    {(void **)&_mega_sync, 256},
	    {(void **)&_mega_amax_val, sizeof(float) * (size_t)sms},
	    {(void **)&_mega_amax_idx, sizeof(int) * (size_t)sms},
	    {(void **)&_mega_tmp, sizeof(float) * c.dim},
	    {(void **)&_mega_hbp, sizeof(float) * ((size_t)k_act * lhid + 512)}, // hb in the w2 tile loop's permuted layout (padded to a chunk)
	    {(void **)&_mega_hb_cnt, sizeof(unsigned) * (size_t)c.n_layers * yalm::mega::kHbMaxSlices},
	    {(void **)&_mega_ll, 8 * (yalm::mega::kLLRep * ((size_t)2 * c.dim + (size_t)k_act * lhid + (size_t)lh * c.head_dim) +
	                              (size_t)sms * lh * c.head_dim + (size_t)sms * lh * 2)},
	};
	size_t total = 0;
	for (auto &s : slots) total += align_up(s.bytes, 256);
	CUDA_CHECK(cudaMalloc(&_arena, total));
	// zeroed on the state's own (non-blocking) stream: the megakernel's barrier words, head counters and flagged
	// hand-over buffers must read 0 at the first launch, which is enqueued on that stream
	CUDA_CHECK(cudaMemsetAsync(_arena, 0, total, _stream));
	size_t off = 0;
	for (auto &s : slots) {
		*s.ptr = static_cast<char *>(_arena) + off;
		off += align_up(s.bytes, 256);
	}
	if (c.tp_size > 1) {
		// tensor parallel: the gathered logits live in this rank's peer-mapped region, where the megakernel's LM
		// head epilogues of all ranks write their rows (one inference state per process uses it)
		const yalm::TpPeerView pv = yalm::tp_peer_view();
		if (pv.ok && pv.world == c.tp_size && c.vocab_size <= yalm::mega::kTpMaxVocab)
			_logits_dev = reinterpret_cast<float *>(pv.base[pv.rank] + yalm::mega::kTpOffLogits);
	}
	CUDA_CHECK(cudaHostAlloc((void **)&_logits, sizeof(float) * c.vocab_size, cudaHostAllocDefault));
	memset(_logits, 0, sizeof(float) * c.vocab_size);
	CUDA_CHECK(cudaStreamSynchronize(_stream)); // the arena is zero before anybody can see its pointers
}

// ------------------------------------------------------------------------------------------------
// Block
// ------------------------------------------------------------------------------------------------
static void *check_tensor(const Tensor *tensor, DType dtype, std::array<int, 4> shape) {
	if (tensor == nullptr) {
		std::cerr << "FATAL: missing tensor" << std::endl;
		abort();
	}
	if (tensor->dtype != dtype || tensor->shape != shape) {
		std::cerr << "FATAL: tensor mismatch for " << tensor->name << std::endl;
		fprintf(stderr, "expected: dtype=%s, shape=[%d,%d,%d,%d]\n", dtype_to_string(dtype).c_str(), shape[0], shape[1],
		        shape[2], shape[3]);
		fprintf(stderr, "got: dtype=%s, shape=[%d,%d,%d,%d]\n", dtype_to_string(tensor->dtype).c_str(),
		        tensor->shape[0], tensor->shape[1], tensor->shape[2], tensor->shape[3]);
		abort();
	}
	return tensor->data;
}

static const Tensor *get_tensor(const YALMData &yalm, const std::string &key) {
	const Tensor *t = yalm.find(key);
	if (!t) {
		std::cerr << "FATAL: missing tensor: " << key << std::endl;
		abort();
	}
	return t;
}

Block::Block(int layer_i, const std::shared_ptr<Config> config, const Tensor *rms_att_weight,
             const Tensor *rms_ffn_weight, const Tensor *wq, const Tensor *wk, const Tensor *wv, const Tensor *wo,
             const Tensor *w1, const Tensor *w2, const Tensor *w3, const Tensor *moegate)
    : _layer_i(layer_i), _config(config) {
	const Config &c = *config;
	const DType dt = c.weight_dtype;
	if (dt != DType::F32 && dt != DType::F16 && dt != DType::F8E4M3) {
		std::cerr << "FATAL: unsupported weight dtype " << dtype_to_string(dt)
		          << " (fp8 checkpoints must be F8_E4M3 + .scale tensors; the reference's unscaled F8_E5M2 has no "
		             "arithmetic defined)"
		          << std::endl;
		abort();
	}
	const int q_dim = c.n_heads * c.head_dim, kv_dim = c.n_kv_heads * c.head_dim;
	_rms_att_weight = static_cast<float *>(check_tensor(rms_att_weight, DType::F32, {c.dim, 0, 0, 0}));
	_rms_ffn_weight = static_cast<float *>(check_tensor(rms_ffn_weight, DType::F32, {c.dim, 0, 0, 0}));
	_wq = check_tensor(wq, dt, {q_dim, c.dim, 0, 0});
	_wk = check_tensor(wk, dt, {kv_dim, c.dim, 0, 0});
	_wv = check_tensor(wv, dt, {kv_dim, c.dim, 0, 0});
	_wo = check_tensor(wo, dt, {c.dim, q_dim, 0, 0});
	if (c.n_experts > 0) {
		_moegate = check_tensor(moegate, dt, {c.n_experts, c.dim, 0, 0});
		_w1 = check_tensor(w1, dt, {c.n_experts, c.hidden_dim, c.dim, 0});
		_w2 = check_tensor(w2, dt, {c.n_experts, c.dim, c.hidden_dim, 0});
		_w3 = check_tensor(w3, dt, {c.n_experts, c.hidden_dim, c.dim, 0});
	} else {
		_w1 = check_tensor(w1, dt, {c.hidden_dim, c.dim, 0, 0});
		_w2 = check_tensor(w2, dt, {c.dim, c.hidden_dim, 0, 0});
		_w3 = check_tensor(w3, dt, {c.hidden_dim, c.dim, 0, 0});
	}
	// The KV cache is created on the device in cuda(); there is no host copy (the reference allocates
	// and zero-fills 2*max_seq_len*kv_dim halfs per layer on the host first, src/model.cpp:171-172).
}

void Block::set_scales(const Tensor *sq, const Tensor *sk, const Tensor *sv, const Tensor *so, const Tensor *s1,
                       const Tensor *s2, const Tensor *s3, const Tensor *sgate) {
	const Config &c = *_config;
	const int q_dim = c.n_heads * c.head_dim, kv_dim = c.n_kv_heads * c.head_dim;
	_sq = static_cast<float *>(check_tensor(sq, DType::F32, {q_dim, 0, 0, 0}));
	_sk = static_cast<float *>(check_tensor(sk, DType::F32, {kv_dim, 0, 0, 0}));
	_sv = static_cast<float *>(check_tensor(sv, DType::F32, {kv_dim, 0, 0, 0}));
	_so = static_cast<float *>(check_tensor(so, DType::F32, {c.dim, 0, 0, 0}));
	if (c.n_experts > 0) {
		_s1 = static_cast<float *>(check_tensor(s1, DType::F32, {c.n_experts, c.hidden_dim, 0, 0}));
		_s2 = static_cast<float *>(check_tensor(s2, DType::F32, {c.n_experts, c.dim, 0, 0}));
		_s3 = static_cast<float *>(check_tensor(s3, DType::F32, {c.n_experts, c.hidden_dim, 0, 0}));
		_sgate = static_cast<float *>(check_tensor(sgate, DType::F32, {c.n_experts, 0, 0, 0}));
	} else {
		_s1 = static_cast<float *>(check_tensor(s1, DType::F32, {c.hidden_dim, 0, 0, 0}));
		_s2 = static_cast<float *>(check_tensor(s2, DType::F32, {c.dim, 0, 0, 0}));
		_s3 = static_cast<float *>(check_tensor(s3, DType::F32, {c.hidden_dim, 0, 0, 0}));
	}
}

Block::~Block() = default; // device memory belongs to the model's arena (or to _own_arena)

// `height` pieces of `width` bytes, piece i read at host + i*spitch + offset -> one dense range of the arena.
static void *upload_strided(yalm::DeviceArena &arena, yalm::Uploader &up, size_t &acct, const void *host, size_t spitch,
                            size_t offset, size_t width, size_t height) {
	if (!host) return nullptr;
	void *dev = arena.take(width * height);
	if (!dev) {
		std::cerr << "FATAL: device arena exhausted (loader sizing bug)" << std::endl;
		abort();
	}
	OP_CHECK(up.put(dev, host, spitch, offset, width, height));
	acct += width * height;
	return dev;
}

// Which host bytes of every tensor this rank keeps.  One list drives both the sizing pass (device_bytes_needed) and the
// upload pass (cuda), so the arena is exactly as large as what goes into it.
struct Block::Piece {
	void **slot;
	const void *host; // captured before any slot is overwritten with its device address
	size_t spitch, offset, width, height;
	bool cont;        // continues the previous piece of the same slot (must land right behind it in the arena)
};

std::vector<Block::Piece> Block::pieces() {
	const Config &c = *_config;
	const size_t B = dtype_size(c.weight_dtype), T = c.tp_size, r = c.tp_rank;
	const size_t q_dim = (size_t)c.n_heads * c.head_dim, kv_dim = (size_t)c.n_kv_heads * c.head_dim;
	const size_t lq = q_dim / T, lkv = kv_dim / T, lhid = (size_t)c.hidden_dim / T, dim = c.dim;
	const size_t E = std::max(1, c.n_experts);
	const yalm::TpDims td{c.dim, c.hidden_dim, c.head_dim, c.n_heads, c.n_kv_heads, c.vocab_size, (int)B};
	std::vector<Piece> out;
	auto whole = [&](void **slot, size_t bytes) { out.push_back({slot, *slot, bytes, 0, bytes, 1, false}); };
	auto slice = [&](void **slot, int tensor, size_t repeat) { // `repeat` stacked experts
		const yalm::TpSlice sl = yalm::tp_slice(td, tensor, c.tp_rank, c.tp_size);
		out.push_back({slot, *slot, sl.spitch, sl.offset, sl.width, sl.height * repeat, false});
	};
	// expert parallelism: whole experts r, r + T, ... of a stacked tensor with `per` bytes per expert, one after the other
	auto experts = [&](void **slot, size_t per) {
		for (size_t e = r, j = 0; e < E; e += T, ++j) out.push_back({slot, *slot, per, e * per, per, 1, j > 0});
	};
	whole((void **)&_rms_att_weight, dim * 4);
	whole((void **)&_rms_ffn_weight, dim * 4);
	// column-parallel (rows of W): q/k/v by head, w1/w3 by hidden unit
	slice(&_wq, yalm::kTpWq, 1);
	slice(&_wk, yalm::kTpWk, 1);
	slice(&_wv, yalm::kTpWv, 1);
	if (c.moe_ep) experts(&_w1, (size_t)c.hidden_dim * dim * B), experts(&_w3, (size_t)c.hidden_dim * dim * B);
	else slice(&_w1, yalm::kTpW1, E), slice(&_w3, yalm::kTpW3, E);
	// row-parallel (columns of W): wo over q_dim, w2 over hidden
	slice(&_wo, yalm::kTpWo, 1);
	if (c.moe_ep) experts(&_w2, (size_t)c.hidden_dim * dim * B);
	else slice(&_w2, yalm::kTpW2, E);
	if (c.n_experts > 0) whole(&_moegate, E * dim * B);
	if (_sq) {
		out.push_back({(void **)&_sq, _sq, lq * 4, r * lq * 4, lq * 4, 1, false});
		out.push_back({(void **)&_sk, _sk, lkv * 4, r * lkv * 4, lkv * 4, 1, false});
		out.push_back({(void **)&_sv, _sv, lkv * 4, r * lkv * 4, lkv * 4, 1, false});
		whole((void **)&_so, dim * 4);
		if (c.moe_ep) {
			experts((void **)&_s1, (size_t)c.hidden_dim * 4), experts((void **)&_s3, (size_t)c.hidden_dim * 4);
			experts((void **)&_s2, dim * 4);
		} else {
			out.push_back({(void **)&_s1, _s1, (size_t)c.hidden_dim * 4, r * lhid * 4, lhid * 4, E, false});
			out.push_back({(void **)&_s3, _s3, (size_t)c.hidden_dim * 4, r * lhid * 4, lhid * 4, E, false});
			whole((void **)&_s2, E * dim * 4);
		}
		if (_sgate) whole((void **)&_sgate, E * 4);
	}
	return out;
}

size_t Block::device_bytes_needed() {
	const Config &c = *_config;
	size_t n = 0;
	for (const Piece &p : pieces()) n += yalm::DeviceArena::padded(p.width * p.height);
	const size_t kv_bytes = (size_t)c.max_seq_len * (c.n_kv_heads / c.tp_size) * c.head_dim * sizeof(f16_t);
	return n + 2 * yalm::DeviceArena::padded(kv_bytes);
}

void Block::cuda(yalm::DeviceArena &arena, yalm::Uploader &up) {
	if (_device != Device::CPU) return;
	_device = Device::CUDA;
	const Config &c = *_config;
	char *prev_end = nullptr;
	for (const Piece &p : pieces()) {
		char *dev = static_cast<char *>(upload_strided(arena, up, _device_bytes, p.host, p.spitch, p.offset, p.width, p.height));
		if (!p.cont) *p.slot = dev;
		else if (dev != prev_end) { // expert blocks are multiples of 256 bytes for every real shape
			std::cerr << "FATAL: expert-parallel load: expert size is not a multiple of the arena alignment" << std::endl;
			abort();
		}
		prev_end = dev + p.width * p.height;
	}
	const size_t kv_bytes = (size_t)c.max_seq_len * c.local_kv_heads() * c.head_dim * sizeof(f16_t);
	_key_cache = static_cast<f16_t *>(arena.take(kv_bytes));
	_value_cache = static_cast<f16_t *>(arena.take(kv_bytes));
	if (!_key_cache || !_value_cache) {
		std::cerr << "FATAL: device arena exhausted (KV cache)" << std::endl;
		abort();
	}
	OP_CHECK(up.zero(_key_cache, kv_bytes));
	OP_CHECK(up.zero(_value_cache, kv_bytes));
	_device_bytes += 2 * kv_bytes;
}

// Stand-alone use of a Block (reference signature, src/model.h:246): its own small arena.
void Block::cuda() {
	if (_device != Device::CPU) return;
	_own_arena = std::make_shared<yalm::DeviceArena>();
	CUDA_CHECK((cudaError_t)_own_arena->reserve(device_bytes_needed()));
	yalm::Uploader up(64u << 20);
	cuda(*_own_arena, up);
	OP_CHECK(up.finish());
}

int Block::enqueue(InferenceState &s) const {
	const Config &c = *_config;
	const int dt = static_cast<int>(c.weight_dtype);
	const int lh = c.local_heads(), lkv = c.local_kv_heads(), lhid = c.local_hidden();
	yalm_stream_t st = s.stream();
	const bool tp = c.tp_size > 1;
	int rc;
	rc = yalm_qkv_rope_append(s.q(), _key_cache, _value_cache, s.x(), _rms_att_weight, _wq, _wk, _wv, _sq, _sk, _sv,
	                          c.dim, lh, lkv, c.head_dim, c.rotary_dim, c.max_seq_len, c.rope_theta, c.norm_eps, c.qkv_clip,
	                          dt, s.ctl(), st);
	if (rc) return rc;
	if (yalm_attn_is_fast_shape(lh, lkv, c.head_dim))
		rc = yalm_attn_decode(s.xb2(), _key_cache, _value_cache, s.q(), s.attn_workspace(), lh, lkv, c.head_dim,
		                      c.max_seq_len, s.ctl(), st);
	else
		rc = yalm_attn_generic(s.xb2(), s.att(), _key_cache, _value_cache, s.q(), lh, lkv, c.head_dim, c.max_seq_len, 0,
		                       s.ctl(), st);
	if (rc) return rc;
	// wo + residual.  Under TP every rank holds a column slice: rank 0 folds the residual into its partial
	// and the all-reduce leaves the new x on every rank.
	if (!tp) {
		rc = yalm_gemv_residual(s.x(), s.x(), s.xb2(), _wo, _so, lh * c.head_dim, c.dim, dt, 1, nullptr, nullptr, st);
		if (rc) return rc;
	} else {
		rc = yalm_gemv_residual(s.xb(), c.tp_rank == 0 ? s.x() : nullptr, s.xb2(), _wo, _so, lh * c.head_dim, c.dim, dt,
		                        1, nullptr, nullptr, st);
		if (rc) return rc;
		rc = yalm::tp_allreduce_sum(s.xb(), s.x(), c.dim, s.stream());
		if (rc) return rc;
	}
	const bool moe = c.n_experts > 0;
	if (s.ffn_timing()) CUDA_CHECK(cudaEventRecordWithFlags(s.ffn_event(_layer_i, 0), s.stream(), cudaEventRecordExternal));
	const int ep_rank = c.moe_ep ? c.tp_rank : 0, ep_size = c.moe_ep ? c.tp_size : 1;
	rc = yalm_ffn_glu_ep(s.hb(), s.x(), _rms_ffn_weight, _w1, _w3, _s1, _s3, _moegate, _sgate, c.dim, lhid,
	                     c.act == ActivationType::SILU ? YALM_ACT_SILU : YALM_ACT_GELU, c.norm_eps, dt, c.n_experts,
	                     c.n_experts_active, s.active_experts(), s.active_experts_weights(), s.moe_weights(), ep_rank, ep_size, st);
	if (rc) return rc;
	if (s.ffn_timing()) CUDA_CHECK(cudaEventRecordWithFlags(s.ffn_event(_layer_i, 1), s.stream(), cudaEventRecordExternal));
	const int k_act = moe ? c.n_experts_active : 1;
	const int *ae = moe ? s.active_experts() : nullptr;
	const float *aw = moe ? s.active_experts_weights() : nullptr;
	if (!tp) {
		rc = yalm_gemv_residual(s.x(), s.x(), s.hb(), _w2, _s2, lhid, c.dim, dt, k_act, ae, aw, st);
	} else {
		rc = yalm_gemv_residual_ep(s.xb(), c.tp_rank == 0 ? s.x() : nullptr, s.hb(), _w2, _s2, lhid, c.dim, dt, k_act, ae, aw, ep_rank,
		                           ep_size, st);
		if (rc) return rc;
		rc = yalm::tp_allreduce_sum(s.xb(), s.x(), c.dim, s.stream());
	}
	return rc;
}

void Block::block(InferenceState &s, int pos, int kv_sink, int kv_pos, int kv_len) const {
	if (_device != Device::CUDA || s.device() != Device::CUDA) {
		std::cerr << "FATAL: yalm_b200 has no CPU execution path; call .cuda() on the model and the state" << std::endl;
		abort();
	}
	int h[YALM_CTL_INTS] = {0, pos, kv_sink, kv_pos, kv_len, 0, 0, 0};
	CUDA_CHECK(cudaMemcpyAsync(s.ctl(), h, sizeof(h), cudaMemcpyHostToDevice, s.stream()));
	CUDA_CHECK(cudaStreamSynchronize(s.stream())); // `h` is a stack buffer
	OP_CHECK(enqueue(s));
}

// ------------------------------------------------------------------------------------------------
// Model
// ------------------------------------------------------------------------------------------------
Model::Model(YALMData &yalm, int context) {
	config = std::make_shared<Config>();
	config->from_yalm(yalm, context);
	const Config &c = *config;
	std::cout << "loading model with dtype: " << dtype_to_string(c.weight_dtype) << std::endl;
	const DType dt = c.weight_dtype;
	const bool fp8 = dt == DType::F8E4M3;
	token_embedding_table = check_tensor(get_tensor(yalm, "model.embed.weight"), dt, {c.vocab_size, c.dim, 0, 0});
	if (fp8)
		embed_scale = static_cast<float *>(
		    check_tensor(get_tensor(yalm, "model.embed.weight.scale"), DType::F32, {c.vocab_size, 0, 0, 0}));
	for (int i = 0; i < c.n_layers; ++i) {
		const std::string p = "model.layers." + std::to_string(i) + ".";
		auto T = [&](const char *n) { return get_tensor(yalm, p + n); };
		auto blk = std::make_shared<Block>(i, config, T("attn.norm.weight"), T("mlp.norm.weight"), T("attn.wq.weight"),
		                                   T("attn.wk.weight"), T("attn.wv.weight"), T("attn.wo.weight"),
		                                   T("mlp.w1.weight"), T("mlp.w2.weight"), T("mlp.w3.weight"),
		                                   c.n_experts > 0 ? T("moegate.weight") : nullptr);
		if (fp8)
			blk->set_scales(T("attn.wq.weight.scale"), T("attn.wk.weight.scale"), T("attn.wv.weight.scale"),
			                T("attn.wo.weight.scale"), T("mlp.w1.weight.scale"), T("mlp.w2.weight.scale"),
			                T("mlp.w3.weight.scale"), c.n_experts > 0 ? T("moegate.weight.scale") : nullptr);
		blocks.emplace_back(blk);
	}
	rms_final_weight = static_cast<float *>(check_tensor(get_tensor(yalm, "model.norm.weight"), DType::F32, {c.dim, 0, 0, 0}));
	_tied = yalm.tensors.count("model.output.weight") == 0; // src/model.cpp:371-377
	if (_tied) {
		wcls = token_embedding_table;
		wcls_scale = embed_scale;
	} else {
		wcls = check_tensor(get_tensor(yalm, "model.output.weight"), dt, {c.vocab_size, c.dim, 0, 0});
		if (fp8)
			wcls_scale = static_cast<float *>(
			    check_tensor(get_tensor(yalm, "model.output.weight.scale"), DType::F32, {c.vocab_size, 0, 0, 0}));
	}
}

Model::~Model() = default; // the arena frees the one device allocation

void Model::cuda() {
	int count = 0;
	if (cudaGetDeviceCount(&count) != cudaSuccess || count == 0) {
		std::cerr << "FATAL: no CUDA device; yalm_b200 has no CPU fallback" << std::endl;
		abort();
	}
	int dev = 0;
	cudaGetDevice(&dev);
	cuda(dev, yalm::tp_rank(), yalm::tp_world());
}

void Model::cuda(int device, int tp_rank, int tp_size) {
	if (_device != Device::CPU) return;
	Config &c = *config;
	const yalm::TpDims td{c.dim, c.hidden_dim, c.head_dim, c.n_heads, c.n_kv_heads, c.vocab_size, (int)dtype_size(c.weight_dtype)};
	if (tp_size < 1 || tp_rank < 0 || tp_rank >= tp_size || !yalm::tp_shards(td, tp_size)) {
		std::cerr << "FATAL: model does not shard " << tp_size << " ways (kv heads " << c.n_kv_heads << ", hidden "
		          << c.hidden_dim << ", vocab " << c.vocab_size << ")" << std::endl;
		abort();
	}
	c.tp_rank = tp_rank, c.tp_size = tp_size;
	// YALM_MOE_EP=1: whole experts per rank instead of a slice of every expert (per-op kernels; the measured alternative)
	const char *ep_env = getenv("YALM_MOE_EP");
	c.moe_ep = ep_env && ep_env[0] == '1' && tp_size > 1 && c.n_experts > 0 && c.n_experts % tp_size == 0;
	_device = Device::CUDA;
	set_cuda_device(device);
	const size_t B = dtype_size(c.weight_dtype), dim = c.dim, V = c.vocab_size, lv = V / tp_size;
	// ---- sizing pass: one device allocation for everything this rank holds ------------------------------------
	using yalm::DeviceArena;
	const yalm::TpSlice cls = yalm::tp_slice(td, yalm::kTpWcls, tp_rank, tp_size);
	size_t need = DeviceArena::padded(V * dim * B) + DeviceArena::padded(dim * 4);
	if (embed_scale) need += DeviceArena::padded(V * 4);
	for (auto &b : blocks) need += b->device_bytes_needed();
	if (!_tied) need += DeviceArena::padded(cls.width * cls.height) + (wcls_scale ? DeviceArena::padded(lv * 4) : 0);
	_arena = std::make_shared<DeviceArena>();
	CUDA_CHECK((cudaError_t)_arena->reserve(need));
	// ---- upload pass: pinned double-buffered staging, slices cut on the host side ---------------------------------
	yalm::Uploader up;
	DeviceArena &ar = *_arena;
	token_embedding_table = upload_strided(ar, up, _device_bytes, token_embedding_table, V * dim * B, 0, V * dim * B, 1);
	if (embed_scale) embed_scale = (float *)upload_strided(ar, up, _device_bytes, embed_scale, V * 4, 0, V * 4, 1);
	for (auto &b : blocks) b->cuda(ar, up);
	for (auto &b : blocks) _device_bytes += b->device_bytes();
	rms_final_weight = (float *)upload_strided(ar, up, _device_bytes, rms_final_weight, dim * 4, 0, dim * 4, 1);
	// LM head: vocabulary rows are split across ranks; tied heads alias the (replicated) embedding table.
	if (_tied) {
		wcls = static_cast<char *>(token_embedding_table) + (size_t)tp_rank * lv * dim * B;
		wcls_scale = embed_scale ? embed_scale + (size_t)tp_rank * lv : nullptr;
	} else {
		wcls = upload_strided(ar, up, _device_bytes, wcls, cls.spitch, cls.offset, cls.width, cls.height);
		if (wcls_scale)
			wcls_scale = (float *)upload_strided(ar, up, _device_bytes, wcls_scale, lv * 4, (size_t)tp_rank * lv * 4, lv * 4, 1);
	}
	OP_CHECK(up.finish());
	_upload_pinned = up.pinned();
	plan_mega();
	CUDA_CHECK(cudaDeviceSynchronize());
}

// ---- megakernel path ---------------------------------------------------------------------------------------
struct Model::MegaState {
	yalm::mega::Plan plan;
	yalm::mega::Layer *layers = nullptr; // device
	void *tmaps = nullptr;               // device: tensor maps of the weight matrices (tensor-core GEMV)
	~MegaState() {
		cudaFree(layers);
		cudaFree(tmaps);
	}
};

void Model::plan_mega() {
	const Config &c = *config;
	const char *env = getenv("YALM_B200_MEGA");
	if (env && env[0] == '0') return;
	// tensor parallel: the in-kernel all-reduce needs the peer-mapped regions (tp.cpp); without them the per-op
	// kernels + NCCL run instead
	const bool tp = c.tp_size > 1;
	if (c.moe_ep) return; // expert parallelism runs on the per-op kernels
	if (tp) {
		const yalm::TpPeerView pv = yalm::tp_peer_view();
		const char *e2 = getenv("YALM_TP_MEGA"); // 0: tensor parallel on the per-op kernels (A/B)
		if ((e2 && e2[0] == '0') || !pv.ok || pv.world != c.tp_size || pv.rank != c.tp_rank || c.tp_size % 2) return;
	}
	auto ms = std::make_shared<MegaState>();
	ms->plan = yalm::mega::plan(static_cast<int>(c.weight_dtype), c.dim, c.local_heads() * c.head_dim,
	                            c.local_kv_heads() * c.head_dim, c.head_dim, c.local_hidden(), c.local_vocab(), c.local_heads(),
	                            c.local_kv_heads(), c.n_experts, c.n_experts_active, c.max_seq_len, c.n_layers, tp, c.vocab_size);
	if (!ms->plan.ok) return;
	std::vector<yalm::mega::Layer> host(blocks.size());
	for (size_t i = 0; i < blocks.size(); ++i) {
		const Block &b = *blocks[i];
		yalm::mega::Layer &L = host[i];
		L.wq = (const char *)b._wq, L.wk = (const char *)b._wk, L.wv = (const char *)b._wv, L.wo = (const char *)b._wo;
		L.w1 = (const char *)b._w1, L.w2 = (const char *)b._w2, L.w3 = (const char *)b._w3, L.gate = (const char *)b._moegate;
		L.rms_att = b._rms_att_weight, L.rms_ffn = b._rms_ffn_weight;
		L.sq = b._sq, L.sk = b._sk, L.sv = b._sv, L.so = b._so, L.s1 = b._s1, L.s2 = b._s2, L.s3 = b._s3, L.sgate = b._sgate;
		L.kc = b._key_cache, L.vc = b._value_cache;
	}
	CUDA_CHECK(cudaMalloc((void **)&ms->layers, sizeof(yalm::mega::Layer) * host.size()));
	CUDA_CHECK(cudaMemcpy(ms->layers, host.data(), sizeof(yalm::mega::Layer) * host.size(), cudaMemcpyHostToDevice));
	if (ms->plan.mma) { // tensor-core GEMV: one tensor map per weight matrix (mega.h: MmaMat order per layer, LM head last)
		using namespace yalm::mega;
		const int q_dim = c.local_heads() * c.head_dim, kv_dim = c.local_kv_heads() * c.head_dim, hid = c.local_hidden();
		std::vector<unsigned char> maps((host.size() * kMmPerLayer + 1) * 128);
		bool ok = true;
		for (size_t i = 0; i < host.size(); ++i) {
			const Layer &L = host[i];
			unsigned char *m = maps.data() + i * kMmPerLayer * 128;
			const int *mt = ms->plan.mma_mt;
			ok = ok && make_weight_map(m + kMmWq * 128, L.wq, q_dim, c.dim, mt[0]) == 0 && make_weight_map(m + kMmWk * 128, L.wk, kv_dim, c.dim, mt[1]) == 0 &&
			     make_weight_map(m + kMmWv * 128, L.wv, kv_dim, c.dim, mt[1]) == 0 && make_weight_map(m + kMmWo * 128, L.wo, c.dim, q_dim, mt[2]) == 0 &&
			     make_weight_map(m + kMmW1 * 128, L.w1, hid, c.dim, mt[3]) == 0 && make_weight_map(m + kMmW3 * 128, L.w3, hid, c.dim, mt[3]) == 0 &&
			     make_weight_map(m + kMmW2 * 128, L.w2, c.dim, hid, mt[4]) == 0;
		}
		ok = ok && make_weight_map(maps.data() + host.size() * kMmPerLayer * 128, wcls, c.local_vocab(), c.dim, ms->plan.mma_mt[5]) == 0;
		if (!ok) {
			std::cerr << "FATAL: tensor maps for the tensor-core GEMV could not be encoded" << std::endl;
			abort();
		}
		CUDA_CHECK(cudaMalloc((void **)&ms->tmaps, maps.size()));
		CUDA_CHECK(cudaMemcpy(ms->tmaps, maps.data(), maps.size(), cudaMemcpyHostToDevice));
	}
	_mega = ms;
	_mega_ok = true;
}

int Model::mega_info(int *out8) const {
	if (!_mega_ok) return YALM_EINVAL;
	const yalm::mega::Plan &p = _mega->plan;
	out8[0] = p.stages, out8[1] = p.vec_floats, out8[2] = p.w2_pipe, out8[3] = p.shape[3][2], out8[4] = p.mma;
	out8[5] = out8[6] = out8[7] = 0;
	return 0;
}

int Model::mega_profile(InferenceState &s, int token, int pos, long long *out, size_t count, int *grid) {
	if (!_mega_ok) return YALM_EINVAL;
	const Config &c = *config;
	const size_t need = (size_t)_mega->plan.grid * c.n_layers * yalm::mega::kProfileSlots;
	if (grid) *grid = _mega->plan.grid;
	if (!out) return 0;
	if (count < need) return YALM_EINVAL;
	long long *dev = nullptr;
	CUDA_CHECK(cudaMalloc((void **)&dev, need * sizeof(long long)));
	CUDA_CHECK(cudaMemsetAsync(dev, 0, need * sizeof(long long), s.stream())); // same stream as the launch below
	// YALM_MEGA_TRACE=<file>: also dump CTA 0's event trace (decode_mega.cuh: trace()) as raw int64
	const char *tf = getenv("YALM_MEGA_TRACE");
	long long *tdev = nullptr;
	const size_t tn = 1 + 2 * (size_t)yalm::mega::kTraceEvents;
	if (tf) {
		CUDA_CHECK(cudaMalloc((void **)&tdev, tn * sizeof(long long)));
		CUDA_CHECK(cudaMemsetAsync(tdev, 0, tn * sizeof(long long), s.stream()));
	}
	_mega_trace = tdev;
	int rc = launch_mega(s, token, pos, yalm::mega::kLogits, dev);
	_mega_trace = nullptr;
	if (rc == 0) rc = (int)cudaStreamSynchronize(s.stream());
	if (rc == 0) rc = (int)cudaMemcpy(out, dev, need * sizeof(long long), cudaMemcpyDeviceToHost);
	if (rc == 0 && tdev) {
		std::vector<long long> h(tn);
		rc = (int)cudaMemcpy(h.data(), tdev, tn * sizeof(long long), cudaMemcpyDeviceToHost);
		if (FILE *f = fopen(tf, "wb")) {
			fwrite(h.data(), sizeof(long long), tn, f);
			fclose(f);
		}
	}
	cudaFree(tdev);
	cudaFree(dev);
	return rc;
}

int Model::launch_mega(InferenceState &s, int token, int pos, int mode, long long *prof) {
	const Config &c = *config;
	yalm::mega::Args a{};
	a.prof = prof;
	a.trace = _mega_trace;
	a.layers = _mega->layers, a.n_layers = c.n_layers;
	a.dim = c.dim, a.q_dim = c.local_heads() * c.head_dim, a.kv_dim = c.local_kv_heads() * c.head_dim, a.head_dim = c.head_dim;
	a.hidden = c.local_hidden(), a.vocab = c.local_vocab(), a.n_heads = c.local_heads(), a.n_kv_heads = c.local_kv_heads();
	a.tp_world = c.tp_size, a.tp_rank = c.tp_rank, a.vocab_base = c.tp_rank * c.local_vocab(), a.x2 = s.xb();
	if (c.tp_size > 1) {
		const yalm::TpPeerView pv = yalm::tp_peer_view();
		if (!pv.ok || pv.world != c.tp_size || s.logits_dev() != reinterpret_cast<float *>(pv.base[pv.rank] + yalm::mega::kTpOffLogits))
			return YALM_EINVAL; // the state must have been created after yalm_tp_init (its logits live in the peer region)
		for (int r = 0; r < pv.world; ++r) a.tp_peer[r] = pv.base[r];
	}
	a.rotary_dim = c.rotary_dim, a.max_seq_len = c.max_seq_len;
	a.act = c.act == ActivationType::SILU ? YALM_ACT_SILU : YALM_ACT_GELU;
	a.n_experts = c.n_experts, a.n_active = c.n_experts_active;
	a.n_sub = c.n_experts > 0 ? c.n_experts_active : 1, a.hb_len = a.n_sub * a.hidden;
	a.theta = c.rope_theta, a.eps = c.norm_eps, a.clip = c.qkv_clip;
	a.embed = (const char *)token_embedding_table, a.embed_scale = embed_scale, a.rms_final = rms_final_weight;
	a.wcls = (const char *)wcls, a.wcls_scale = wcls_scale;
	a.x = s.x(), a.q = s.q(), a.att_out = s.xb2(), a.hb = s.hb(), a.logits = s.logits_dev();
	a.moe_logits = s.moe_weights(), a.active_w = s.active_experts_weights(), a.active_e = s.active_experts();
	a.ws_acc = s.mega_acc(), a.ws_ml = s.mega_ml(), a.head_counters = s.mega_counters(), a.sync = s.mega_sync();
	a.amax_val = s.mega_amax_val(), a.amax_idx = s.mega_amax_idx();
	constexpr size_t rep = yalm::mega::kLLRep;
	a.ll_x = s.mega_ll(), a.ll_hb = a.ll_x + rep * 2 * c.dim, a.ll_att = a.ll_hb + rep * (size_t)a.hb_len;
	a.ll_acc = a.ll_att + rep * a.q_dim, a.ll_ml = a.ll_acc + (size_t)_mega->plan.grid * a.q_dim;
	// launch number behind the hand-over flags: per state on one GPU; under tensor parallelism the peers' mailboxes belong
	// to the process, so the counter does too (every rank issues the same launches in the same order)
	static unsigned tp_seq = 0;
	a.seq = c.tp_size > 1 ? ++tp_seq : s.next_mega_seq();
	a.w2_tmp = s.mega_tmp();
	a.hbp = s.mega_hbp(), a.hb_cnt = s.mega_hb_cnt();
	a.tmaps = _mega->tmaps;
	a.ctl = s.ctl(), a.tokens_out = s.tokens_dev();
	a.token = token, a.pos = pos, a.mode = mode;
	return yalm::mega::launch(_mega->plan, a, s.stream());
}

int Model::launches_per_forward(InferenceMode mode) const {
	if (_mega_ok) return 1;
	// embed + 5 kernels per block (qkv+rope+append, flash-decode, wo+res, ffn-glu, w2+res) [+ lm head]
	// (+ 2 NCCL all-reduce kernels per block and one all-gather under tensor parallelism)
	const int per_block = 5 + (config->tp_size > 1 ? 2 : 0);
	int n = 1 + per_block * config->n_layers;
	if (mode == InferenceMode::OUTPUT_LOGITS) n += 1 + (config->tp_size > 1 ? 1 : 0);
	return n;
}

// Enqueue (eagerly or into a capturing stream) everything after the embedding.
int Model::enqueue_forward(InferenceState &s, InferenceMode mode, bool /*from_host*/, int tail) {
	const Config &c = *config;
	const int dt = static_cast<int>(c.weight_dtype);
	yalm_stream_t st = s.stream();
	int rc;
	const bool greedy_tail = tail == kTailGreedy;
	if (tail != kTailHost) { // token/pos come from the device-side control block
		rc = yalm_embed(s.x(), token_embedding_table, embed_scale, c.dim, dt, s.ctl(), -1, 0, c.max_seq_len, st);
		if (rc) return rc;
	}
	for (auto &b : blocks) {
		rc = b->enqueue(s);
		if (rc) return rc;
	}
	if (mode == InferenceMode::HYDRATE_KV_CACHE) return 0;
	const int lv = c.local_vocab();
	rc = yalm_lm_head(s.logits_dev() + (size_t)c.tp_rank * lv, s.x(), rms_final_weight, wcls, wcls_scale, c.dim, lv,
	                  c.norm_eps, dt, st);
	if (rc) return rc;
	if (c.tp_size > 1) {
		rc = yalm::tp_allgather(s.logits_dev() + (size_t)c.tp_rank * lv, s.logits_dev(), lv, s.stream());
		if (rc) return rc;
	}
	if (greedy_tail) return yalm_argmax_advance(s.logits_dev(), c.vocab_size, s.ctl(), c.max_seq_len, s.tokens_dev(), st);
	if (tail == kTailDevice || !s.host_logits()) return 0; // the caller samples on the device (yalm_sample_advance / Sampler::*_device)
	return (int)cudaMemcpyAsync(s.logits(), s.logits_dev(), sizeof(float) * c.vocab_size, cudaMemcpyDeviceToHost,
	                            s.stream());
}

static void capture_graph(CudaGraph &g, cudaStream_t st, const std::function<int()> &body);

void Model::forward(InferenceState &s, int token, int pos, InferenceMode mode) {
	if (s.device() != _device) {
		std::cerr << "FATAL: inference state device mismatch" << std::endl;
		abort();
	}
	if (_device != Device::CUDA) {
		std::cerr << "FATAL: yalm_b200 has no CPU execution path; call model.cuda() and state.cuda() first" << std::endl;
		abort();
	}
	const Config &c = *config;
	if (token < 0 || token >= c.vocab_size || pos < 0) {
		std::cerr << "FATAL: token/pos out of range" << std::endl;
		abort();
	}
	s.set_mode(mode);
	if (_mega_ok) {
		// one persistent kernel per token; token / pos travel as kernel arguments
		OP_CHECK(launch_mega(s, token, pos,
		                     mode == InferenceMode::OUTPUT_LOGITS ? yalm::mega::kLogits : yalm::mega::kHydrate));
		if (mode == InferenceMode::OUTPUT_LOGITS) {
			if (s.host_logits())
				CUDA_CHECK(cudaMemcpyAsync(s.logits(), s.logits_dev(), sizeof(float) * c.vocab_size, cudaMemcpyDeviceToHost,
				                           s.stream()));
			CUDA_CHECK(cudaStreamSynchronize(s.stream()));
		}
		CUDA_CHECK(cudaGetLastError());
		return;
	}
	CudaGraph &g = s.host_logits() || mode == InferenceMode::HYDRATE_KV_CACHE ? s.graph() : s.output_dev_graph();
	// The embedding is an ordinary launch: token and pos travel as kernel arguments (copied at launch
	// time), everything after it reads positions from the device control block and is one graph replay.
	OP_CHECK(yalm_embed(s.x(), token_embedding_table, embed_scale, c.dim, static_cast<int>(c.weight_dtype), s.ctl(),
	                    token, pos, c.max_seq_len, s.stream()));
	if (!g.is_created) {
		// First call in this mode: run eagerly (this also sizes every kernel's shared memory and
		// occupancy), then capture the same sequence for all later calls.
		OP_CHECK(enqueue_forward(s, mode, true, kTailHost));
		CUDA_CHECK(cudaStreamSynchronize(s.stream()));
		capture_graph(g, s.stream(), [&]() { return enqueue_forward(s, mode, true, kTailHost); });
	} else {
		g.launch(s.stream());
	}
	if (mode == InferenceMode::OUTPUT_LOGITS) CUDA_CHECK(cudaStreamSynchronize(s.stream()));
	CUDA_CHECK(cudaGetLastError());
}

static void capture_graph(CudaGraph &g, cudaStream_t st, const std::function<int()> &body) {
	CUDA_CHECK(cudaStreamBeginCapture(st, cudaStreamCaptureModeThreadLocal));
	int rc = body();
	cudaError_t end = cudaStreamEndCapture(st, &g.graph);
	if (rc != 0 || end != cudaSuccess) {
		fprintf(stderr, "yalm_b200: graph capture failed (%d, %s)\n", rc, cudaGetErrorString(end));
		abort();
	}
	CUDA_CHECK(cudaGraphInstantiate(&g.instance, g.graph, 0));
	g.is_created = true;
}

int Model::generate_greedy(InferenceState &s, int token, int pos, int n_gen, int *tokens_out, float *logits_out,
                           float *device_ms) {
	return generate(s, token, pos, n_gen, 0.f, 0ull, tokens_out, logits_out, device_ms);
}

// n_gen decode steps without host round trips.  temperature == 0: greedy (first-max arg-max inside the megakernel, or
// yalm_argmax_advance behind the per-op graph); > 0: Sampler::sample on the device (yalm_sample_advance) with
// r = uniform(seed, step).  Only the token ids (4 bytes each) come back.
int Model::generate(InferenceState &s, int token, int pos, int n_gen, float temperature, unsigned long long seed,
                    int *tokens_out, float *logits_out, float *device_ms) {
	if (_device != Device::CUDA || s.device() != Device::CUDA) return YALM_ENOGPU;
	const Config &c = *config;
	if (n_gen < 1 || n_gen > s.tokens_capacity() || token < 0 || token >= c.vocab_size || pos < 0 || !tokens_out ||
	    !(temperature >= 0.f))
		return YALM_EINVAL;
	const bool greedy = temperature == 0.f;
	cudaStream_t st = s.stream();
	// seed the device control block: token, pos, ring position, step = 0
	const int kv_sink = pos >= c.max_seq_len ? KV_SINKS : 0;
	int h[YALM_CTL_INTS] = {token, pos, kv_sink, kv_sink + (pos - kv_sink) % (c.max_seq_len - kv_sink),
	                        pos >= c.max_seq_len ? c.max_seq_len : pos + 1, 0, 0, 0};
	CUDA_CHECK(cudaMemcpyAsync(s.ctl(), h, sizeof(h), cudaMemcpyHostToDevice, st));
	CUDA_CHECK(cudaStreamSynchronize(st));
	cudaEvent_t e0, e1;
	CUDA_CHECK(cudaEventCreate(&e0));
	CUDA_CHECK(cudaEventCreate(&e1));
	auto finish = [&](int done) -> int {
		CUDA_CHECK(cudaEventRecord(e1, st));
		CUDA_CHECK(cudaMemcpyAsync(tokens_out, s.tokens_dev(), sizeof(int) * n_gen, cudaMemcpyDeviceToHost, st));
		CUDA_CHECK(cudaStreamSynchronize(st));
		if (device_ms) {
			CUDA_CHECK(cudaEventElapsedTime(device_ms, e0, e1));
			if (done && n_gen > done) *device_ms = *device_ms * n_gen / (n_gen - done); // scale to n_gen steps when step 0 ran eagerly
		}
		cudaEventDestroy(e0);
		cudaEventDestroy(e1);
		return (int)cudaGetLastError();
	};
	auto sample = [&]() { return yalm_sample_advance(s.logits_dev(), c.vocab_size, s.ctl(), c.max_seq_len, s.tokens_dev(), temperature, seed, st); };
	auto keep_logits = [&](int i) {
		if (logits_out)
			CUDA_CHECK(cudaMemcpyAsync(logits_out + (size_t)i * c.vocab_size, s.logits_dev(), sizeof(float) * c.vocab_size,
			                           cudaMemcpyDeviceToHost, st));
	};
	if (_mega_ok) {
		CUDA_CHECK(cudaEventRecord(e0, st));
		for (int i = 0; i < n_gen; ++i) {
			// token / pos from the device control block; greedy: the arg-max and the advance happen inside the kernel
			int rc = launch_mega(s, -1, 0, greedy ? yalm::mega::kGreedy : yalm::mega::kLogits);
			if (rc) return rc;
			keep_logits(i);
			if (!greedy && (rc = sample())) return rc;
		}
		return finish(0);
	}
	CudaGraph &g = greedy ? s.greedy_graph() : s.device_graph();
	const int tail = greedy ? kTailGreedy : kTailDevice;
	int done = 0;
	if (!g.is_created) {
		int rc = enqueue_forward(s, InferenceMode::OUTPUT_LOGITS, false, tail);
		if (rc) return rc;
		keep_logits(0);
		if (!greedy && (rc = sample())) return rc;
		CUDA_CHECK(cudaStreamSynchronize(st));
		capture_graph(g, st, [&]() { return enqueue_forward(s, InferenceMode::OUTPUT_LOGITS, false, tail); });
		done = 1;
	}
	CUDA_CHECK(cudaEventRecord(e0, st));
	for (int i = done; i < n_gen; ++i) {
		g.launch(st);
		keep_logits(i);
		if (!greedy) {
			int rc = sample();
			if (rc) return rc;
		}
	}
	return finish(done);
}

int Model::profile_ffn(InferenceState &s, int token, int pos, int n_steps, double *mean_us) {
	if (_device != Device::CUDA || s.device() != Device::CUDA) return YALM_ENOGPU;
	const Config &c = *config;
	if (n_steps < 1 || !mean_us || token < 0 || token >= c.vocab_size || pos < 0) return YALM_EINVAL;
	if (_mega_ok) return YALM_EINVAL; // no separate w1/w3 kernel on the megakernel path
	cudaStream_t st = s.stream();
	const int kv_sink = pos >= c.max_seq_len ? KV_SINKS : 0;
	int h[YALM_CTL_INTS] = {token, pos, kv_sink, kv_sink + (pos - kv_sink) % (c.max_seq_len - kv_sink),
	                        pos >= c.max_seq_len ? c.max_seq_len : pos + 1, 0, 0, 0};
	CUDA_CHECK(cudaMemcpyAsync(s.ctl(), h, sizeof(h), cudaMemcpyHostToDevice, st));
	CUDA_CHECK(cudaStreamSynchronize(st));
	s.enable_ffn_timing(true);
	CudaGraph &g = s.profile_graph();
	if (!g.is_created) capture_graph(g, st, [&]() { return enqueue_forward(s, InferenceMode::OUTPUT_LOGITS, false, kTailGreedy); });
	s.enable_ffn_timing(false);
	double total = 0;
	for (int i = 0; i < n_steps; ++i) {
		g.launch(st);
		CUDA_CHECK(cudaStreamSynchronize(st));
		for (int l = 0; l < c.n_layers; ++l) {
			float ms = 0.f;
			CUDA_CHECK(cudaEventElapsedTime(&ms, s.ffn_event(l, 0), s.ffn_event(l, 1)));
			total += ms;
		}
	}
	*mean_us = 1000.0 * total / ((double)n_steps * c.n_layers);
	return 0;
}
