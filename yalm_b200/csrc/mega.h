// Whole-token persistent decode kernel ("megakernel"): one cooperative launch per token, one CTA per SM.
// A producer warp streams every weight / KV tile of the forward pass through a shared-memory ring with
// 1-D bulk async copies (TMA engine, mbarrier completion) in a fixed per-CTA order, running ahead of the
// consumer warps by the depth of the ring, so the HBM pipe stays full across the ~160 phase
// boundaries of a 32-layer model instead of draining at every kernel launch.  Phases hand their vectors over as
// flagged 64-bit words (Args::ll_*); one grid barrier per layer remains.
// Host-side planning and launch; the kernel is in decode_mega.cu.
#pragma once

#include <cuda_runtime_api.h>
#include <stddef.h>
#include <stdint.h>

namespace yalm {
namespace mega {

constexpr int kStageBytes = 8192;  // one ring stage = one weight tile or one K+V tile pair
constexpr int kConsumerWarps = 11;
constexpr int kProducerWarps = 1; // all 32 lanes issue: lane l takes tile l of each batch of min(stages, 32) tiles
constexpr int kThreads = (kConsumerWarps + kProducerWarps) * 32;
constexpr int kMaxStages = 24;
constexpr int kSmemHeader = 2048; // (a multiple of 1024: the ring stages are 1024-byte aligned for swizzled tensor copies) // rope table, reduction scratch, release counters, mbarriers in front of the ring
// Pipelined w1/w3 -> w2 hand-over (dense models, Args::w2_pipe; decode_mega.cuh "pipelined hb"): hb travels as plain fp32
// in the permuted shared-memory layout, sliced by the K splits of w2; each slice has a completion counter and is pulled
// into one of kHbBufs shared-memory buffers by a bulk copy; w2 walks its tiles K-split-major, so a warp carries the
// partial sums of all its row groups (at most kHbMaxGroups) across the slices.
constexpr int kHbMaxSlices = 16, kHbBufs = 4, kHbMaxGroups = 3;
// Tensor-core GEMV (fp16 dense models, Args::mma; decode_mega.cuh "tensor-core GEMV"): a ring stage holds 16-row tiles of
// 256 columns fetched with ONE 3-D tensor copy (64 columns x MT rows x 4 column blocks, SWIZZLE_128B) and is multiplied
// with mma.sync.m16n8k16 against x kept as an fp16 hi + lo pair; the K stages of a tile are spread over the consumer warps
// and summed in warp order through kMmaSlots partial slots in shared memory.
constexpr int kMmaKCols = 256, kMmaSlots = 8;
enum MmaMat { kMmWq = 0, kMmWk, kMmWv, kMmWo, kMmW1, kMmW2, kMmW3, kMmPerLayer }; // tensor map l * kMmPerLayer + i; the LM head's is last
constexpr int kTraceEvents = 1 << 16;
// Mixture of experts (Mixtral: 8 experts, 2 active; the reference has MoE on the CPU only, src/infer.cpp:100-132, 347-384;
// its CUDA path asserts, src/infer.cu:865-867).  The router runs in the staging step of the w1/w3 phase (every CTA
// recomputes the n_experts logits and the top-k, bit-identically), the producer waits for it and streams the rows of the
// SELECTED experts (expert-indirect bases), and w2 runs as n_active sub-phases — one per selected expert, same row -> CTA
// -> warp -> lane map, the running sum x + w_0 W2[e_0] hb_0 (+ ...) waits in Args::w2_tmp with the lane that made it —
// so the summation order is the reference's (src/infer.cpp:377-383).  Dense models: n_sub = 1.
constexpr int kMegaMaxExperts = 16, kMegaMaxActive = 4;
#ifndef YALM_MEGA_LL_REP
#define YALM_MEGA_LL_REP 4
#endif
constexpr int kFlagPhaseBits = 12;  // hand-over flag = launch number << kFlagPhaseBits | (phase + 1), see decode_mega.cuh
constexpr int kLLRep = YALM_MEGA_LL_REP; // copies of every flagged hand-over vector (decode_mega.cuh); 1, 2 or 4
constexpr int kProfileSlots = 6 * 8; // per (CTA, layer): phases x slots, see decode_mega.cuh

// ---- tensor parallelism inside the kernel (N GPUs of one NVSwitch domain) ---------------------------------------------
// Every rank owns one region of device memory that all peers map (CUDA IPC, tp.cpp).  The row -> CTA -> warp map of a
// phase is the same on every rank, so the all-reduce behind wo / w2 is an exchange between the warps that own the same
// rows: the epilogue pushes the warp's partial rows into mailbox [set][rank] of EVERY rank's region over NVLink as
// flagged 64-bit words {fp32 value, flag} (the hand-over protocol of decode_mega.cuh; NCCL's LL protocol relies on the
// same 8-byte atomicity), and at the end of the phase the warp polls the world's words for ITS rows in its own region,
// adds them to the residual in rank order — the same bits on every rank — and publishes the finished rows as local
// flagged words, exactly as on a single GPU.  No cross-GPU barrier, no system-scope fence, no atomics on the data path;
// the all-reduce has no kernel of its own.  The LM head scatters its logits rows and arg-max candidates with plain
// stores; ONE cross-GPU barrier per token (at the very end) makes them visible.
constexpr int kTpMaxWorld = 8;
constexpr int kTpCapFloats = 8192;  // largest dim
constexpr int kTpMaxGrid = 256;     // CTAs per rank (arg-max candidates)
constexpr int kTpMaxVocab = 1 << 17;
constexpr size_t kTpOffGen = 0;     // u32: cross-GPU barriers completed by earlier launches (same on every rank)
constexpr size_t kTpOffBar = 128;   // u32: arrival counter, monotonic (wrap-safe compare)
constexpr size_t kTpOffMail = 1024;  // u64 [2 sets][kTpMaxWorld][kTpCapFloats]: flagged partial rows; set 0 behind wo, 1 behind w2
constexpr size_t kTpOffAmaxV = kTpOffMail + 2ull * kTpMaxWorld * kTpCapFloats * 8; // float [world][kTpMaxGrid]
constexpr size_t kTpOffAmaxI = kTpOffAmaxV + (size_t)kTpMaxWorld * kTpMaxGrid * 4;   // int   [world][kTpMaxGrid]
constexpr size_t kTpOffLogits = kTpOffAmaxI + (size_t)kTpMaxWorld * kTpMaxGrid * 4;  // float [vocab], all ranks' rows
constexpr size_t kTpRegionBytes = kTpOffLogits + (size_t)kTpMaxVocab * 4;

// Per-layer device pointers (weights are (out,in) row-major; KV cache is head-major, see yalm_b200.h).
struct Layer {
	const char *wq, *wk, *wv, *wo, *w1, *w2, *w3, *gate;
	const float *rms_att, *rms_ffn;
	const float *sq, *sk, *sv, *so, *s1, *s2, *s3, *sgate; // fp8 per-row scales (null otherwise)
	uint16_t *kc, *vc;
};

enum Mode { kHydrate = 0, kLogits = 1, kGreedy = 2 };

struct Args {
	const Layer *layers;
	int n_layers;
	int dim, q_dim, kv_dim, head_dim, hidden, vocab, n_heads, n_kv_heads, rotary_dim, max_seq_len, act;
	int n_experts, n_active;
	int n_sub;  // w2 sub-phases per layer: 1 (dense) or n_active (MoE)
	int hb_len; // n_sub * hidden: hb holds act(w1 x) * (w3 x) of every selected expert
	float theta, eps, clip;
	const char *embed;
	const float *embed_scale, *rms_final;
	const char *wcls;
	const float *wcls_scale;
	// activations (global, fp32)
	float *x, *q, *att_out, *hb, *logits;
	float *moe_logits, *active_w;
	int *active_e;
	// attention split partials: acc (n_kv, grid, G, HD), ml (n_kv, grid, G, 2)
	float *ws_acc, *ws_ml;
	unsigned *head_counters; // (n_kv) arrival counters, zero between uses
	unsigned *sync;          // [0] grid-barrier counter, [1] exit counter; zero between launches
	float *amax_val;         // (grid) per-CTA arg-max candidates
	int *amax_idx;
	int *ctl, *tokens_out;
	int token, pos; // token >= 0: taken from here; < 0: from ctl (device greedy loop)
	int mode;
	int stages, vec_floats;
	int shape[5][4]; // per GEMV kind (qkv, wo, w1/w3, w2, lm head): rows per tile, columns per tile, K splits, row groups
	int l2_ahead;    // L2 look-ahead of the producer in tiles per CTA (0: off)
	int debug;       // timing experiments only (YALM_MEGA_DEBUG): 1 skip GEMV math, 4 skip attention math, 8 no copies
	long long *trace; // optional event trace of CTA 0: [0] count, then kTraceEvents (tag, clock) pairs; null = off
	long long *prof; // optional phase profile, (grid, n_layers, 6 phases, 8 slots) clock64 values; null = off
	// tensor parallelism (tp_world > 1: every dimension above is this rank's local share, dim excepted)
	int tp_world, tp_rank;
	int vocab_base;             // first vocabulary row of this rank
	float *x2;                  // second residual buffer (x ping-pongs across the in-kernel all-reduces)
	// Flagged hand-over buffers (single GPU, see decode_mega.cuh "flagged hand-over"): every entry is one 64-bit word
	// {fp32 value, u32 flag}, flag = seq << kFlagPhaseBits | (producing phase + 1); zero before the first launch.
	unsigned long long *ll_x;   // [2][kLLRep][dim]: [0] x after wo, [1] x after w2
	unsigned long long *ll_hb;  // [kLLRep][hb_len]
	unsigned long long *ll_att; // [kLLRep][q_dim]
	unsigned long long *ll_acc; // attention partials, (n_kv, grid, G, HD)
	unsigned long long *ll_ml;  // (n_kv, grid, G, 2): {max, flag}, {sum, flag}
	unsigned seq;               // launch number of this inference state, starts at 1
	float *w2_tmp;              // [dim]: MoE, x + the earlier experts' contributions, private to the lane that wrote it
	// pipelined hb (w2_pipe != 0): hbp = hb in the permuted layout (xs_slot) the w2 tile loop reads; hb_cnt = completion
	// counters [n_layers][kHbMaxSlices] (zero between launches); hb_slice = floats per slice (= columns per w2 tile);
	// hb_off = float offset of the slice buffers inside the vec region (behind the phases' own input vector)
	float *hbp;
	unsigned *hb_cnt;
	int w2_pipe, hb_slice, hb_off;
	// tensor-core GEMV (mma != 0): 128-byte CUtensorMap records (MmaMat order per layer, LM head last); box rows per matrix
	// class: [0] wq [1] wk / wv [2] wo [3] w1 / w3 (rows of each) [4] w2 [5] LM head
	const void *tmaps;
	int mma, mma_mt[6];
	char *tp_peer[kTpMaxWorld]; // every rank's region (kTpOff*), [tp_rank] = this rank's own
};

struct Plan {
	bool ok = false;
	int dtype = 0, grid = 0, stages = 0, vec_floats = 0;
	bool tp = false;
	int w2_pipe = 0, hb_slice = 0, hb_off = 0; // pipelined hb (Args)
	int mma = 0, mma_mt[6] = {}; // tensor-core GEMV (Args)
	size_t smem = 0;
	const void *kernel = nullptr;
	int shape[5][4] = {};
};

// Can this model shape run on the megakernel on the current device?  Under tensor parallelism (`tp`) the dimensions
// are the rank's local shares (q_dim, kv_dim, hidden, vocab, heads) and `full_vocab` the unsharded vocabulary.
Plan plan(int dtype, int dim, int q_dim, int kv_dim, int head_dim, int hidden, int vocab, int n_heads,
          int n_kv_heads, int n_experts, int n_active, int max_seq_len, int n_layers, bool tp = false, int full_vocab = 0);
// Bytes of the attention partial workspace (acc + ml) for `grid` CTAs.
size_t attn_ws_floats(int grid, int n_heads, int head_dim);
// One token.  Returns a cudaError_t.
int launch(const Plan &p, const Args &a, cudaStream_t st);
// Tensor-core GEMV: tensor map (128 bytes at `map128`) of a (rows, K) row-major fp16 matrix on the device for boxes of
// 64 columns x box_rows rows x 4 column blocks, 128-byte swizzle, zero fill past the last row.  Returns 0 or YALM_EINVAL.
int make_weight_map(void *map128, const void *dev_ptr, int rows, int K, int box_rows);
// Failure record of the kernel's timeouts (decode_mega.cuh: die): 8 words of pinned, device-mapped host memory, all zero
// while nothing has gone wrong.  describe_fault() formats it ("" if clean); fault_timeout_cycles(): YALM_MEGA_TIMEOUT_S.
unsigned *fault_record();
long long fault_timeout_cycles();
const char *describe_fault();

} // namespace mega
} // namespace yalm
