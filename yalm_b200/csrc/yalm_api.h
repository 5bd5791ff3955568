// C++ surface of the B200-native forward pass.  Type, method and field names follow the reference's
// src/model.h (Config :41-68, InferenceState :84-190, Block :193-301, Model :303-327, test entry points
// :353-384) so that yalm's main.cpp / test.cpp / sampler.cpp-style callers compile against it; what is
// underneath is different: every op is an sm_100a kernel reached through the C ABI in
// include/yalm_b200.h, positions live in device memory, and one CUDA graph per mode is captured once
// and replayed with no per-token node updates.
//
// There is no CPU execution path: Model::forward on Device::CPU reports the error and aborts.
#pragma once

#include "codec.h"

#include <cuda_runtime_api.h>

#include <memory>
#include <string>
#include <vector>

#include "../../include/yalm_b200.h"

namespace yalm {
class DeviceArena;
class Uploader;
} // namespace yalm

constexpr int KV_SINKS = 2; // src/model.h:12

enum class ActivationType { GELU, SILU };
enum class LayerNormType { RMSNorm };
enum class Device { CPU, CUDA };
enum class InferenceMode {
	HYDRATE_KV_CACHE, // only append to the KV cache
	OUTPUT_LOGITS     // also leave next-token logits in InferenceState::logits()
};

struct Config {
	int dim = 0, hidden_dim = 0, head_dim = 0, n_layers = 0, n_heads = 0, n_kv_heads = 0, vocab_size = 0,
	    max_seq_len = 0;
	float rope_theta = 10000.f;
	int rotary_dim = 0;
	float norm_eps = 1e-5f;
	ActivationType act = ActivationType::GELU;
	LayerNormType norm_type = LayerNormType::RMSNorm;
	float qkv_clip = 3.402823466e+38f;
	int n_experts = 0, n_experts_active = 0;
	DType weight_dtype = DType::F32;
	// tensor-parallel shard this process holds (net-new; 0 of 1 = the whole model)
	int tp_rank = 0, tp_size = 1;
	// expert parallelism for the MoE FFN (net-new, the measured alternative of SURVEY.md 8(e)): rank r holds whole experts
	// e with e % tp_size == r instead of a 1 / tp_size slice of every expert; attention and the LM head stay tensor-parallel
	bool moe_ep = false;

	// If nonzero `context` is supplied, max sequence length is set to `context` (src/model.cpp:31-36).
	void from_yalm(YALMData &yalm, int context = 0);
	// Exact weight + KV bytes one token reads at position `pos`; unlike the
	// reference's (src/model.cpp:77-102) the LM head is counted at its real element size.
	size_t active_bytes(size_t pos) const;
	// per-shard sizes
	int local_heads() const { return n_heads / tp_size; }
	int local_kv_heads() const { return n_kv_heads / tp_size; }
	int local_hidden() const { return moe_ep ? hidden_dim : hidden_dim / tp_size; }
	int local_experts() const { return moe_ep ? n_experts / tp_size : n_experts; }
	int local_vocab() const { return vocab_size / tp_size; }
};

// One captured forward pass.
struct CudaGraph {
	cudaGraph_t graph = nullptr;
	cudaGraphExec_t instance = nullptr;
	bool is_created = false;
	void launch(cudaStream_t s);
	void destroy();
};

struct Model;

// All per-token scratch.  Host buffers until cuda() is called, device buffers afterwards — except
// logits(), which is always host memory (pinned once on the device) so Sampler-style callers keep working.
struct InferenceState {
	explicit InferenceState(const std::shared_ptr<Config> config);
	~InferenceState();
	InferenceState(const InferenceState &) = delete;

	float *x() const { return _x; }
	float *xb() const { return _xb; }
	float *xb(int head) const { return _xb + _config->head_dim * head; }
	float *xb2() const { return _xb2; }
	float *xb2(int head) const { return _xb2 + _config->head_dim * head; }
	float *hb() const { return _hb; }
	float *hb2() const { return _hb2; }
	float *q() const { return _q; }
	float *q(int head) const { return _q + _config->head_dim * head; }
	float *k() const { return _k; }
	float *v() const { return _v; }
	float *att() const { return _att; }
	float *att(int head) const { return _att + _config->max_seq_len * head; }
	float *moe_weights() const { return _moe_weights; }
	float *active_experts_weights() const { return _active_experts_weights; }
	int *active_experts() const { return _active_experts; }
	float *logits() const { return _logits; }

	void cuda();
	Device device() const { return _device; }
	cudaStream_t stream() const { return _stream; }
	InferenceMode mode() const { return _mode; }
	void set_mode(InferenceMode mode) { _mode = mode; }
	CudaGraph &graph() { return _mode == InferenceMode::HYDRATE_KV_CACHE ? _hydrate_graph : _output_graph; }

	// device-side decode control (include/yalm_b200.h: YALM_CTL_INTS)
	int *ctl() const { return _ctl; }
	void *attn_workspace() const { return _attn_ws; }
	float *logits_dev() const { return _logits_dev; }
	int *tokens_dev() const { return _tokens_dev; }
	int tokens_capacity() const { return _tokens_cap; }
	CudaGraph &greedy_graph() { return _greedy_graph; }
	CudaGraph &output_dev_graph() { return _output_dev_graph; } // forward(OUTPUT_LOGITS) with host_logits() off
	CudaGraph &device_graph() { return _device_graph; } // forward pass whose logits stay on the device (device-side sampling)
	// false: OUTPUT_LOGITS forwards leave the logits on the device only (logits_dev()); the caller samples there
	// (Sampler::sample_device / sample_prob_device) and 4 bytes per token cross PCIe instead of vocab * 4
	void set_host_logits(bool on) { _host_logits = on; }
	bool host_logits() const { return _host_logits; }
	const Config &config() const { return *_config; }
	// Optional kernel timing: when enabled, Block::enqueue brackets the w1/w3 GLU kernel (the dominant
	// kernel: 54 % of a layer's weight bytes) with an event pair per layer.
	void enable_ffn_timing(bool on);
	bool ffn_timing() const { return _ffn_timing; }
	cudaEvent_t ffn_event(int layer, int which) const { return _ffn_events[2 * layer + which]; }
	CudaGraph &profile_graph() { return _profile_graph; }
	// megakernel workspace (decode_mega.cu): attention partials, head arrival counters, grid-barrier words,
	// per-CTA arg-max candidates
	float *mega_acc() const { return _mega_acc; }
	float *mega_ml() const { return _mega_ml; }
	unsigned *mega_counters() const { return _mega_counters; }
	unsigned *mega_sync() const { return _mega_sync; }
	float *mega_amax_val() const { return _mega_amax_val; }
	int *mega_amax_idx() const { return _mega_amax_idx; }
	unsigned long long *mega_ll() const { return _mega_ll; }
	float *mega_tmp() const { return _mega_tmp; }
	float *mega_hbp() const { return _mega_hbp; }
	unsigned *mega_hb_cnt() const { return _mega_hb_cnt; }
	unsigned next_mega_seq() { return ++_mega_seq; }
	void set_mega_seq(unsigned v) { _mega_seq = v; } // tests: start the hand-over flags' launch counter near its wrap

private:
	std::shared_ptr<Config> _config;
	Device _device = Device::CPU;
	cudaStream_t _stream = nullptr;
	InferenceMode _mode = InferenceMode::OUTPUT_LOGITS;
	CudaGraph _hydrate_graph, _output_graph, _greedy_graph, _profile_graph, _device_graph, _output_dev_graph;
	bool _host_logits = true;
	bool _ffn_timing = false;
	std::vector<cudaEvent_t> _ffn_events;

	float *_x = nullptr, *_xb = nullptr, *_xb2 = nullptr, *_hb = nullptr, *_hb2 = nullptr;
	float *_q = nullptr, *_k = nullptr, *_v = nullptr, *_att = nullptr;
	float *_moe_weights = nullptr, *_active_experts_weights = nullptr;
	int *_active_experts = nullptr;
	float *_logits = nullptr;     // host (vocab)
	float *_logits_dev = nullptr; // device (vocab)
	int *_ctl = nullptr;
	void *_attn_ws = nullptr;
	int *_tokens_dev = nullptr;
	int _tokens_cap = 0;
	float *_mega_acc = nullptr, *_mega_ml = nullptr, *_mega_amax_val = nullptr;
	unsigned *_mega_counters = nullptr, *_mega_sync = nullptr;
	int *_mega_amax_idx = nullptr;
	float *_mega_tmp = nullptr; // x + first column half of w2 (YALM_MEGA_W2_SPLIT)
	float *_mega_hbp = nullptr;        // pipelined hb (mega.h: Args::hbp, hb_cnt)
	unsigned *_mega_hb_cnt = nullptr;
	unsigned long long *_mega_ll = nullptr; // flagged hand-over buffers: x[2][dim] | hb | att | acc | ml (mega.h: Args::ll_*)
	unsigned _mega_seq = 0;
	void *_arena = nullptr; // one device allocation behind all of the above
};

/* Transformer block */
struct Block {
	Block(int layer_i, const std::shared_ptr<Config> config, const Tensor *rms_att_weight, const Tensor *rms_ffn_weight,
	      const Tensor *wq, const Tensor *wk, const Tensor *wv, const Tensor *wo, const Tensor *w1, const Tensor *w2,
	      const Tensor *w3, const Tensor *moegate);
	~Block();
	Block(const Block &) = delete;

	float *rms_att_weight() const { return _rms_att_weight; }
	float *rms_ffn_weight() const { return _rms_ffn_weight; }
	template <typename T> T *wq() const { return static_cast<T *>(_wq); }
	template <typename T> T *wk() const { return static_cast<T *>(_wk); }
	template <typename T> T *wv() const { return static_cast<T *>(_wv); }
	template <typename T> T *wo() const { return static_cast<T *>(_wo); }
	template <typename T> T *w1() const { return static_cast<T *>(_w1); }
	template <typename T> T *w2() const { return static_cast<T *>(_w2); }
	template <typename T> T *w3() const { return static_cast<T *>(_w3); }
	template <typename T> T *moegate() const { return static_cast<T *>(_moegate); }
	f16_t *key_cache() const { return _key_cache; }
	f16_t *value_cache() const { return _value_cache; }

	// Forward this block with explicit ring coordinates (reference signature, src/model.h:246-254):
	// writes them to s.ctl() on s.stream() and enqueues the block's kernels.
	void block(InferenceState &s, int pos, int kv_sink, int kv_pos, int kv_len) const;
	// Enqueue the block's kernels, coordinates taken from s.ctl() on the device.
	int enqueue(InferenceState &s) const;
	// Upload this block's (shard of the) weights and create its KV cache: stand-alone (own arena), or into the
	// model's arena through the model's pinned staging stream (csrc/loader.h).
	void cuda();
	void cuda(yalm::DeviceArena &arena, yalm::Uploader &up);
	size_t device_bytes_needed(); // arena bytes cuda() will take (padded)

	// fp8 per-row scales (null unless weight_dtype == F8E4M3)
	void set_scales(const Tensor *sq, const Tensor *sk, const Tensor *sv, const Tensor *so, const Tensor *s1,
	                const Tensor *s2, const Tensor *s3, const Tensor *sgate);
	size_t device_bytes() const { return _device_bytes; }

private:
	friend struct Model;
	int _layer_i = 0;
	std::shared_ptr<Config> _config;
	Device _device = Device::CPU;
	float *_rms_att_weight = nullptr, *_rms_ffn_weight = nullptr;
	void *_wq = nullptr, *_wk = nullptr, *_wv = nullptr, *_wo = nullptr;
	void *_w1 = nullptr, *_w2 = nullptr, *_w3 = nullptr, *_moegate = nullptr;
	float *_sq = nullptr, *_sk = nullptr, *_sv = nullptr, *_so = nullptr, *_s1 = nullptr, *_s2 = nullptr, *_s3 = nullptr,
	      *_sgate = nullptr;
	f16_t *_key_cache = nullptr, *_value_cache = nullptr;
	std::shared_ptr<yalm::DeviceArena> _own_arena; // stand-alone cuda() only
	size_t _device_bytes = 0;
	struct Piece;
	std::vector<Piece> pieces();
};

struct Model {
	std::shared_ptr<Config> config;
	std::vector<std::shared_ptr<Block>> blocks;
	void *token_embedding_table = nullptr; // (vocab_size, dim)
	float *embed_scale = nullptr;          // fp8: (vocab_size)
	float *rms_final_weight = nullptr;     // (dim)
	void *wcls = nullptr;                  // (vocab_size / tp, dim)
	float *wcls_scale = nullptr;

	explicit Model(YALMData &yalm, int context = 0);
	~Model();
	Model(const Model &) = delete;

	void forward(InferenceState &s, int token, int pos, InferenceMode mode = InferenceMode::OUTPUT_LOGITS);
	// Upload to the current shard's GPU.  device < 0: cudaSetDevice is left alone except for the
	// reference default of device 0 when nothing was selected (src/model.cpp:385-386).
	void cuda();
	void cuda(int device, int tp_rank, int tp_size);
	Device device() const { return _device; }

	// n_gen greedy steps without host round trips: forward + on-device first-max arg-max + advance.
	// Returns 0 or an error code; see yalm_generate_greedy in include/yalm_b200.h.
	int generate_greedy(InferenceState &s, int token, int pos, int n_gen, int *tokens_out, float *logits_out,
	                    float *device_ms);
	// Same loop with Sampler::sample (src/sampler.cpp:40-65) on the device: temperature 0 = greedy, > 0 = draw with
	// r = uniform(seed, step).  See yalm_generate in include/yalm_b200.h.
	int generate(InferenceState &s, int token, int pos, int n_gen, float temperature, unsigned long long seed,
	             int *tokens_out, float *logits_out, float *device_ms);
	// n_steps greedy steps with an event pair around every layer's w1/w3 GLU kernel; returns the mean
	// duration of that kernel in microseconds (CUDA events on the launching stream).
	int profile_ffn(InferenceState &s, int token, int pos, int n_steps, double *mean_us);
	// Batched prefill / perplexity (net-new; the reference scores and hydrates token by token,
	// src/main.cpp:91-97, 174-184): n consecutive positions pos0 .. pos0 + n - 1 in one pass — dense tcgen05 GEMMs over the
	// chunk, causal attention over the KV ring, bulk KV append.  On return the KV cache holds the chunk, logits() /
	// logits_dev() hold the LAST position's logits, and, if targets != null, probs_out[i] = softmax(logits_i)[targets[i]]
	// (Sampler::sample_prob per position).  Returns 0 or an error code; supports_batch() says whether this model /
	// shard shape has the path (dense fp16 weights, unsharded), batch_capacity() the largest n per call.
	bool supports_batch() const;
	int batch_capacity() const;
	int forward_batch(InferenceState &s, const int *tokens, int n, int pos0, const int *targets, float *probs_out);
	size_t device_bytes() const { return _device_bytes; }
	// loader facts for tests / bench: number of device allocations behind the model (1) and whether the upload went through pinned staging
	int device_allocations() const { return _arena ? 1 : 0; }
	bool upload_was_pinned() const { return _upload_pinned; }
	int launches_per_forward(InferenceMode mode) const;
	// true when forward() runs as ONE persistent cooperative kernel per token (decode_mega.cu) instead of
	// the per-op kernel graph: unsharded dense models whose shapes the tile planner accepts.  Set
	// YALM_B200_MEGA=0 in the environment before cuda() to force the per-op path.
	bool uses_megakernel() const { return _mega_ok; }
	int mega_info(int *out8) const; // {ring stages, vec floats, pipelined hb, K splits of w2, tensor-core GEMV, 0, 0, 0}; YALM_EINVAL without the megakernel
	// One forward pass with the kernel's phase profile on: clock64 stamps per (CTA, layer, phase, slot), see
	// decode_mega.cuh.  out == null: only *grid is returned.
	int mega_profile(InferenceState &s, int token, int pos, long long *out, size_t count, int *grid);

private:
	enum { kTailHost = 0, kTailGreedy = 1, kTailDevice = 2 }; // what follows the LM head: D2H copy of the logits | device arg-max + advance | nothing
	int enqueue_forward(InferenceState &s, InferenceMode mode, bool from_host, int tail);
	int launch_mega(InferenceState &s, int token, int pos, int mode, long long *prof = nullptr);
	void plan_mega();
	bool _mega_ok = false;
	long long *_mega_trace = nullptr; // debug: event trace buffer for the next launch
	struct MegaState;
	std::shared_ptr<MegaState> _mega;
	struct PrefillState; // chunk workspace of forward_batch (prefill.cu), created on first use
	std::shared_ptr<PrefillState> _prefill;
	Device _device = Device::CPU;
	bool _tied = false;
	std::shared_ptr<yalm::DeviceArena> _arena; // ONE device allocation: every weight shard + every KV cache of this rank
	bool _upload_pinned = false;
	size_t _device_bytes = 0;
};

////////////////////////////////////////
// Exposed for tests (src/model.h:353-384).  Host pointers in and out; the *_cuda functions run the
// sm_100a kernels.  The *_cpu twins of the reference are NOT part of this library: the CPU path is
// the oracle and lives under oracle/.
////////////////////////////////////////
void mha_cuda(float *xout, float *att, f16_t *kb, f16_t *vb, float *q, int head_dim, int kv_len, int max_seq_len,
              int n_heads, int n_kv_heads);
template <typename T> void matmul_cuda(float *xout, float *x, T *w, int n, int d);
template <typename T>
void ffn_cuda(float *xout, float *x, T *w1, T *w2, T *w3, int hidden_dim, int dim, ActivationType act);
