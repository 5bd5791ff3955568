"""-m gpu: the whole-token persistent kernel (decode_mega.cuh) — one cooperative launch per token with the
weight / KV tiles streamed through a shared-memory TMA ring — against the oracle and against the per-op kernel
path (YALM_B200_MEGA=0) on the same models: tile shapes with and without K splits, every compiled
(head_dim, group) pair, sliding window + sinks, on-device greedy loop."""
import os

import numpy as np
import pytest

from conftest import LOGIT_TOL, golden_case
from yalm_b200 import api
from yalm_b200 import yalm_file as yf

pytestmark = pytest.mark.gpu


def run(img, prompt, n_gen, context=0, mega=True, device_loop=False, pipe=None, mma=None):
    """pipe: None = the planner's choice, True / False = YALM_MEGA_W2_PIPE=1 / 0 (pipelined hb slices forced on / off);
    mma likewise for YALM_MEGA_MMA (GEMV phases on the tensor cores)."""
    env = {"YALM_B200_MEGA": "1" if mega else "0"}
    if pipe is not None:
        env["YALM_MEGA_W2_PIPE"] = "1" if pipe else "0"
    if mma is not None:
        env["YALM_MEGA_MMA"] = "1" if mma else "0"
    old = {k: os.environ.get(k) for k in env}
    os.environ.update(env)
    try:
        m = api.Model(img, context).cuda()
    finally:
        for k, v in old.items():
            if v is None:
                del os.environ[k]
            else:
                os.environ[k] = v
    used = m.uses_megakernel()
    if pipe is not None and used:
        assert m.mega_info()["w2_pipe"] == pipe, m.mega_info()
    if mma is not None and used:
        assert m.mega_info()["mma"] == mma, m.mega_info()
    s = api.InferenceState(m)
    if device_loop:
        for pos, tok in enumerate(prompt[:-1]):
            s.forward(tok, pos, api.MODE_HYDRATE)
        toks, lg = s.generate_greedy(prompt[-1], len(prompt) - 1, n_gen, keep_logits=True)
    else:
        toks, lg = s.generate(prompt, n_gen, keep_logits=True)
    s.close()
    m.close()
    return toks, lg, used


def test_path_selection():
    _, img = golden_case("tiny-fp16")
    assert run(img, [1, 2], 2)[2] is True
    assert run(img, [1, 2], 2, mega=False)[2] is False
    _, moe = golden_case("tiny-moe-fp16")
    assert run(moe, [1, 2], 2)[2] is True           # MoE: router + expert-indirect rows inside the whole-token kernel
    assert run(moe, [1, 2], 2, mega=False)[2] is False


@pytest.mark.parametrize("name", ["tiny-fp32", "tiny-fp16", "tiny-gelu-clip-fp32", "tiny-gelu-clip-fp16", "tiny-moe-fp32", "tiny-moe-fp16",
                                  "tiny-kv8-fp16"])
def test_mega_matches_reference_golden_and_per_op_path(name):
    fx, img = golden_case(name)
    prompt, n = fx["prompt"].tolist(), len(fx["tokens"])
    tm, lm, used = run(img, prompt, n)
    assert used
    assert np.abs(lm[fx["keep"]] - fx["logits"]).max() < LOGIT_TOL
    assert tm == fx["tokens"].tolist()
    to, lo, _ = run(img, prompt, n, mega=False)
    assert np.abs(lm - lo).max() < 2e-3
    assert tm == to
    td, ld, _ = run(img, prompt, n, device_loop=True)      # on-device arg-max + control-block advance
    assert td == tm
    assert np.abs(ld - lm).max() < 1e-5


CUSTOM = {
    # K-split tiles: fp32 rows of hidden 9216 (36 KB) exceed one 32 KB ring stage
    "ksplit-fp32": (yf.Arch(256, 9216, 32, 1, 8, 2, 512, 64, 10000.0, name="ksplit"), "fp32"),
    # head_dim 64 / group 4 (Llama-3.2-1B geometry), ragged vocab tile count
    "hd64-fp16": (yf.Arch(512, 1024, 64, 2, 8, 2, 1000, 256, 500000.0, name="hd64"), "fp16"),
    # head_dim 128 / group 8 (70B geometry)
    "hd128g8-fp16": (yf.Arch(1024, 1536, 128, 1, 8, 1, 512, 128, 500000.0, name="hd128g8"), "fp16"),
    "hd128g4-fp8": (yf.Arch(1024, 2048, 128, 2, 8, 2, 512, 128, 1000000.0, name="hd128g4"), "fp8"),
}


@pytest.mark.parametrize("case", sorted(CUSTOM))
def test_mega_shapes_vs_oracle(oracle, case):
    arch, dt = CUSTOM[case]
    img = bytes(yf.random_model(arch, dt, seed=3))
    ref_img = bytes(yf.dequantized_fp32_image(img)) if dt == "fp8" else img
    prompt = [1, 5, 9, 200, 77]
    to, lo = oracle.OracleModel(ref_img).generate(prompt, 40, keep_logits=True)
    tg, lg, used = run(img, prompt, 40)
    assert used
    assert np.abs(lg - lo).max() < LOGIT_TOL
    assert tg == to


def test_mega_sliding_window_and_sinks(oracle):
    """-T 24 on a 2-layer model, 90 steps: the ring wraps 3 times, the two sink keys are re-rotated every step,
    and the producer's prefetched KV tiles must not leak stale sink / current rows into the scores."""
    arch = yf.Arch(256, 512, 32, 2, 8, 2, 512, 64, 10000.0, name="tiny-win")
    img = bytes(yf.random_model(arch, "fp16", seed=21))
    to, lo = oracle.OracleModel(img, context=24).generate([1, 2, 3], 90, keep_logits=True)
    tg, lg, used = run(img, [1, 2, 3], 90, context=24)
    assert used
    assert np.abs(lg - lo).max() < LOGIT_TOL
    assert tg == to


def test_mega_long_context_many_tiles(oracle):
    """kv_len up to 700 at head_dim 32 (256-row KV tiles, several per head) and units spread over all CTAs."""
    arch = yf.Arch(256, 512, 32, 1, 8, 2, 512, 1024, 10000.0, name="tiny-long")
    img = bytes(yf.random_model(arch, "fp16", seed=8))
    prompt = [(i * 37 + 1) % 512 for i in range(690)]
    to, lo = oracle.OracleModel(img).generate(prompt, 10, keep_logits=True)
    tg, lg, used = run(img, prompt, 10)
    assert used
    assert np.abs(lg - lo).max() < LOGIT_TOL
    assert tg == to


def test_mega_attention_at_depth_head_dim_128(oracle):
    """Mistral attention geometry (head_dim 128, 4 query heads per KV head) at kv_len 3000+: 190 16-row KV tiles per
    head spread over all CTAs, split partials merged by the last CTA of each head — the whole-token kernel's own
    attention phase, not the per-op flash-decode kernel."""
    arch = yf.Arch(1024, 2048, 128, 2, 8, 2, 512, 4096, 1000000.0, name="deep-hd128")
    img = bytes(yf.random_model(arch, "fp16", seed=5))
    prompt = [(i * 131 + 7) % 509 + 3 for i in range(3050)]
    to, lo = oracle.OracleModel(img).generate(prompt, 12, keep_logits=True)
    tg, lg, used = run(img, prompt, 12)
    assert used
    assert np.abs(lg - lo).max() < LOGIT_TOL
    assert tg == to


def test_mega_fp8_window_1024_two_wraps(oracle):
    """fp8 weights, -T 1024, 2300 positions: the ring wraps twice and the two sink keys are re-rotated ~1300 times
    (src/infer.cpp:303-317, 483-485); oracle = the reference fp32 arithmetic on the de-quantised weights."""
    arch = yf.Arch(1024, 2048, 128, 2, 8, 2, 512, 4096, 1000000.0, name="win-hd128")
    img = bytes(yf.random_model(arch, "fp8", seed=9))
    ref_img = bytes(yf.dequantized_fp32_image(img))
    prompt = [(i * 57 + 5) % 509 + 3 for i in range(2290)]
    to, lo = oracle.OracleModel(ref_img, context=1024).generate(prompt, 10, keep_logits=True)
    tg, lg, used = run(img, prompt, 10, context=1024)
    assert used
    assert np.abs(lg - lo).max() < LOGIT_TOL
    assert tg == to


def test_mega_mistral_geometry_8_layers(oracle):
    """The headline model's exact layer shape (dim 4096, hidden 14336, 32 / 8 heads x 128, vocab 32000) on 8 layers:
    every tile shape, K split and rotation the 32-layer bench uses."""
    arch = yf.slim(yf.ARCHS["mistral-7b"], 8)
    img = memoryview(yf.random_model_streamed(arch, "fp16", seed=0, device="cuda"))     # 4 GB: drawn on the GPU
    prompt = [1, 5, 9, 200, 77, 3100, 31999]
    to, lo = oracle.OracleModel(img).generate(prompt, 10, keep_logits=True)
    tg, lg, used = run(img, prompt, 10)
    assert used
    assert np.abs(lg - lo).max() < LOGIT_TOL
    assert tg == to
    td, _, _ = run(img, prompt, 10, device_loop=True)
    assert td == to


def test_mega_launch_counter_wrap():
    """The hand-over flags carry the state's launch number in 32 - kFlagPhaseBits = 20 bits: seed it just below the wrap
    and decode across it — same tokens, same logits bits as a fresh state."""
    fx, img = golden_case("tiny-fp16")
    m = api.Model(img, 0).cuda()
    s = api.InferenceState(m)
    first_t, first_l = s.generate_greedy(1, 0, 40, keep_logits=True)
    s.close()
    s = api.InferenceState(m)
    s.set_launch_seq((1 << 20) - 17)
    t, l = s.generate_greedy(1, 0, 40, keep_logits=True)
    assert t == first_t
    assert (l == first_l).all()
    s.close()
    m.close()


def test_mega_deterministic():
    fx, img = golden_case("tiny-fp16")
    a = run(img, fx["prompt"].tolist(), 30)[1]
    b = run(img, fx["prompt"].tolist(), 30)[1]
    assert (a == b).all()


def test_mega_flagged_handover_two_states_interleaved():
    """The phase hand-over words carry (launch number of the state, phase) flags and live in per-state buffers: a second
    inference state of the same model driven in between — repeating every step in hydrate and in device-greedy mode; the
    KV cache belongs to the model (src/model.h:193-301), so it rewrites the same rows with the same bits — must not
    change a single bit of the first state's run (a stale word of the other state, an earlier launch or an earlier
    phase accepted as current would show up here)."""
    fx, img = golden_case("tiny-fp16")
    prompt = fx["prompt"].tolist()
    m = api.Model(img, 0).cuda()
    assert m.uses_megakernel()
    ref_state = api.InferenceState(m)
    ref_t, ref_l = ref_state.generate(prompt, 12, keep_logits=True)
    ref_state.close()
    sa, sb = api.InferenceState(m), api.InferenceState(m)

    def step(tok, pos, mode):
        sa.forward(tok, pos, mode)
        sb.forward(tok, pos, api.MODE_HYDRATE)
        sb.generate_greedy(tok, pos, 1)

    for pos in range(len(prompt)):
        step(prompt[pos], pos, api.MODE_LOGITS if pos + 1 == len(prompt) else api.MODE_HYDRATE)
    toks, lgs = [], []
    for i in range(12):
        lg = sa.logits().copy()
        toks.append(int(np.argmax(lg)))
        lgs.append(lg)
        if i + 1 < 12:
            step(toks[-1], len(prompt) + i, api.MODE_LOGITS)
    assert toks == ref_t
    assert (np.stack(lgs) == ref_l).all()
    sa.close()
    sb.close()
    m.close()


def test_mega_many_launches_stay_deterministic():
    """600 launches on one state (the flag's launch number keeps growing): the same 40 tokens at the start and at the end."""
    fx, img = golden_case("tiny-fp16")
    m = api.Model(img, 0).cuda()
    s = api.InferenceState(m)
    first = s.generate_greedy(1, 0, 40)
    for _ in range(13):
        s.generate_greedy(1, 0, 40)
    assert s.generate_greedy(1, 0, 40) == first
    s.close()
    m.close()


def test_stream_probe_runs():
    us, by = api.bench_op("stream", iters=3, warmup=0)
    assert by / us / 1e3 > 1000.0        # GB/s: anything on a B200 streams faster than 1 TB/s


# Pipelined hb (decode_mega.cuh "pipelined hb"): w2 K-split-major with hb arriving slice by slice through bulk copies.  The
# planner switches it on only where it buys ring stages (70B-class shapes); here it is forced on (YALM_MEGA_W2_PIPE=1).
PIPE = {
    # 3 slices, 1-2 row groups per CTA
    "fp16-3slices": (yf.Arch(512, 6144, 64, 3, 8, 2, 1000, 256, 500000.0, name="pipe16"), "fp16", 24),
    # 9 slices of 1024 floats, fewer row groups than CTAs (some CTAs own no w2 rows but walk every slice)
    "fp32-9slices": (yf.Arch(256, 9216, 32, 2, 8, 2, 512, 64, 10000.0, name="pipe32"), "fp32", 24),
    # fp8 rows with scales, 2 slices of 4096
    "fp8-2slices": (yf.Arch(1024, 8192, 128, 2, 8, 2, 512, 128, 1000000.0, name="pipe8"), "fp8", 24),
    # 70B-like width: 28 row groups per CTA = 3 per warp (the register-rotated group loop), gelu
    "fp16-wide-3groups": (yf.Arch(8192, 4096, 128, 1, 64, 8, 512, 64, 500000.0, act_type="gelu", name="pipewide"), "fp16", 6),
    # the headline geometry (7 slices of 2048, 13-14 row groups per CTA = 2 per warp), two layers
    "mistral-2layers": (yf.slim(yf.ARCHS["mistral-7b"], 2), "fp16", 8),
}


@pytest.mark.parametrize("case", sorted(PIPE))
def test_mega_pipelined_hb_vs_oracle_and_flagged_path(oracle, case):
    arch, dt, n = PIPE[case]
    img = bytes(yf.random_model(arch, dt, seed=11))
    ref_img = bytes(yf.dequantized_fp32_image(img)) if dt == "fp8" else img
    prompt = [1, 5, 9, 200, 77]
    to, lo = oracle.OracleModel(ref_img).generate(prompt, n, keep_logits=True)
    tp_, lp, used = run(img, prompt, n, pipe=True)
    assert used
    assert np.abs(lp - lo).max() < LOGIT_TOL
    assert tp_ == to
    tf, lf, _ = run(img, prompt, n, pipe=False)          # same kernel, hb as flagged words: same rows, same order of sums
    assert tf == tp_
    assert np.abs(lf - lp).max() < 1e-4
    td, ld, _ = run(img, prompt, n, pipe=True, device_loop=True)
    assert td == tp_
    assert np.abs(ld - lp).max() < 1e-5


def test_mega_pipelined_hb_window_wrap_and_many_launches(oracle):
    """Pipelined hb with a sliding window (-T 32, 150 steps: the slice counters are zeroed by every launch's last CTA and
    reused by the next) — tokens and logits against the oracle."""
    arch = yf.Arch(512, 6144, 64, 2, 8, 2, 1000, 256, 500000.0, name="pipewin")
    img = bytes(yf.random_model(arch, "fp16", seed=12))
    to, lo = oracle.OracleModel(img, context=32).generate([1, 2, 3], 150, keep_logits=True)
    tg, lg, used = run(img, [1, 2, 3], 150, context=32, pipe=True)
    assert used
    assert np.abs(lg - lo).max() < LOGIT_TOL
    assert tg == to


# Tensor-core GEMV (decode_mega.cuh "tensor-core GEMV"): 16-row weight tiles through 3-D tensor copies, mma.sync against x as
# an fp16 hi + lo pair.  Forced on (YALM_MEGA_MMA=1) and off on the same image.
MMA = {
    # 2 K stages per tile (fewer than warps: most warps park zero partials), head_dim 64
    "small-k512": (yf.Arch(512, 1024, 64, 3, 8, 2, 1000, 256, 500000.0, name="mma512"), 24),
    # gelu + clip, head_dim 128, 4 K stages
    "gelu-clip-k1024": (yf.Arch(1024, 2048, 128, 2, 8, 2, 512, 128, 1000000.0, act_type="gelu", qkv_clip=1.5, name="mma1024"), 24),
    # 70B-like width: 56 rows per CTA in w2 / wo (4 tiles of 14), 32 K stages, head group 8
    "wide-8192": (yf.Arch(8192, 4096, 128, 1, 64, 8, 512, 64, 500000.0, name="mmawide"), 6),
    # the headline geometry, two layers
    "mistral-2layers": (yf.slim(yf.ARCHS["mistral-7b"], 2), 8),
}


@pytest.mark.parametrize("case", sorted(MMA))
def test_mega_tensor_core_gemv_vs_oracle_and_simt_path(oracle, case):
    arch, n = MMA[case]
    img = bytes(yf.random_model(arch, "fp16", seed=13))
    prompt = [1, 5, 9, 200, 77]
    to, lo = oracle.OracleModel(img).generate(prompt, n, keep_logits=True)
    tm, lm, used = run(img, prompt, n, mma=True)
    assert used
    assert np.abs(lm - lo).max() < LOGIT_TOL
    assert tm == to
    ts, ls, _ = run(img, prompt, n, mma=False)
    assert ts == tm
    assert np.abs(ls - lm).max() < 2e-3                   # x = hi + lo to 2^-22, sums in another order
    td, ld, _ = run(img, prompt, n, mma=True, device_loop=True)
    assert td == tm
    assert np.abs(ld - lm).max() < 1e-5


def test_mega_tensor_core_gemv_window_wrap(oracle):
    """Tensor-core GEMV with a sliding window (-T 32, 150 steps, sinks re-rotated) — tokens and logits against the oracle,
    and bit-identical logits on a second state (the partial sums are added in warp order)."""
    arch = yf.Arch(512, 1024, 64, 2, 8, 2, 1000, 256, 500000.0, name="mmawin")
    img = bytes(yf.random_model(arch, "fp16", seed=15))
    to, lo = oracle.OracleModel(img, context=32).generate([1, 2, 3], 150, keep_logits=True)
    tg, lg, used = run(img, [1, 2, 3], 150, context=32, mma=True)
    assert used
    assert np.abs(lg - lo).max() < LOGIT_TOL
    assert tg == to
    tg2, lg2, _ = run(img, [1, 2, 3], 150, context=32, mma=True)
    assert (lg2 == lg).all()
