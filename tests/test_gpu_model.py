"""-m gpu: whole-model parity through Model::forward (C ABI handles): greedy tokens identical to the
reference's golden sequences and logits within the north-star tolerance; fp8 against the fp32 path on
de-quantised weights; host-driven forward vs on-device greedy loop; linearity-style invariants at size."""
import numpy as np
import pytest

from conftest import GOLDEN_CASES, LOGIT_TOL, golden_case
from yalm_b200 import api
from yalm_b200 import yalm_file as yf

pytestmark = pytest.mark.gpu


def run_gpu(img, prompt, n_gen, context=0, device_loop=False):
    m = api.Model(img, context).cuda()
    s = api.InferenceState(m)
    if not device_loop:
        toks, lg = s.generate(prompt, n_gen, keep_logits=True)
    else:
        for pos, tok in enumerate(prompt[:-1]):
            s.forward(tok, pos, api.MODE_HYDRATE)
        toks, lg = s.generate_greedy(prompt[-1], len(prompt) - 1, n_gen, keep_logits=True)
    s.close()
    m.close()
    return toks, lg


@pytest.mark.parametrize("name", GOLDEN_CASES)
def test_greedy_and_logits_match_reference_golden(name):
    fx, img = golden_case(name)
    toks, lg = run_gpu(img, fx["prompt"].tolist(), len(fx["tokens"]))
    err = np.abs(lg[fx["keep"]] - fx["logits"]).max()
    assert err < LOGIT_TOL, f"max |dlogit| {err}"
    assert toks == fx["tokens"].tolist()


@pytest.mark.parametrize("name", ["tiny-fp16", "tiny-moe-fp16"])
def test_device_greedy_loop_equals_host_loop(name):
    """yalm_generate_greedy (on-device first-max argmax, no host round trip) == forward()+host argmax."""
    fx, img = golden_case(name)
    toks, lg = run_gpu(img, fx["prompt"].tolist(), len(fx["tokens"]), device_loop=True)
    assert toks == fx["tokens"].tolist()
    assert np.abs(lg[fx["keep"]] - fx["logits"]).max() < LOGIT_TOL


@pytest.mark.parametrize("arch", ["tiny", "tiny-moe"])
def test_fp8_matches_fp32_path_on_dequantised_weights(oracle, arch):
    """fp8 has no reference arithmetic: oracle = CPU fp32 path on the de-quantised model."""
    img8 = bytes(yf.random_model(yf.ARCHS[arch], "fp8", seed=14))
    img32 = bytes(yf.dequantized_fp32_image(img8))
    o = oracle.OracleModel(img32)
    prompt = [1, 17, 300, 42, 5]
    to, lo = o.generate(prompt, 70, keep_logits=True)
    tg, lg = run_gpu(img8, prompt, 70)
    assert np.abs(lg - lo).max() < LOGIT_TOL
    assert tg == to


def test_sliding_window_context_flag(oracle):
    """-T n replaces max_seq_len (src/model.cpp:31-36): window 16 on the tiny model, 60 steps past it."""
    img = bytes(yf.random_model(yf.ARCHS["tiny"], "fp16", seed=14))
    o = oracle.OracleModel(img, context=16)
    to, lo = o.generate([1, 2, 3], 60, keep_logits=True)
    tg, lg = run_gpu(img, [1, 2, 3], 60, context=16)
    assert np.abs(lg - lo).max() < LOGIT_TOL
    assert tg == to


def test_mistral_shaped_layers_vs_oracle(oracle):
    """Production head/hidden geometry (dim 4096, hidden 14336, 32 q / 8 kv heads, head_dim 128, vocab 32000),
    2 layers so the CPU oracle finishes in seconds; exercises the 128-dim flash-decode fast path in-model."""
    arch = yf.slim(yf.ARCHS["mistral-7b"], 2)
    img = bytes(yf.random_model(arch, "fp16", seed=5))
    o = oracle.OracleModel(img)
    prompt = [1, 100, 2000, 31999]
    to, lo = o.generate(prompt, 12, keep_logits=True)
    tg, lg = run_gpu(img, prompt, 12)
    assert np.abs(lg - lo).max() < LOGIT_TOL
    assert tg == to


def test_determinism_and_state_reuse():
    """Two independent states over one model give bit-identical logits; a state replays identically."""
    fx, img = golden_case("tiny-fp16")
    m = api.Model(img).cuda()
    outs = []
    for _ in range(2):
        s = api.InferenceState(m)
        _, lg = s.generate(fx["prompt"].tolist(), 20, keep_logits=True)
        outs.append(lg)
        s.close()
    m.close()
    assert (outs[0] == outs[1]).all()


def test_loader_is_one_arena_through_pinned_staging(tmp_path):
    """the reference does one cudaMalloc + one pageable copy per tensor (src/model.cpp:185-211); here a
    model is one device allocation filled through pinned double-buffered staging — from memory and from an mmap'd file,
    with a staging buffer smaller than the tensors (YALM_LOAD_STAGING_MB) so that pieces span several buffers."""
    import os
    fx, img = golden_case("tiny-moe-fp16")
    path = tmp_path / "m.yalm"
    path.write_bytes(img)
    os.environ["YALM_LOAD_STAGING_MB"] = "1"
    try:
        for src in (img, str(path)):
            m = api.Model(src).cuda()
            assert m.load_info() == {"device_allocations": 1, "pinned_staging": True}
            s = api.InferenceState(m)
            toks = s.generate(fx["prompt"].tolist(), 16)
            assert toks == fx["tokens"].tolist()[:16]
            s.close()
            m.close()
    finally:
        del os.environ["YALM_LOAD_STAGING_MB"]
