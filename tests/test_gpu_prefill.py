"""-m gpu: batched prefill / perplexity — the tcgen05 GEMM against numpy, and Model::forward_batch
against its oracle, the reference's token loop (src/main.cpp:91-97, 174-184): same per-position next-token probabilities,
same last-position logits (max |delta| <= 2e-2, the north-star tolerance), same KV rows (fp16, +-1 ulp: the GEMM sums in a
different order), same greedy continuation, perplexity equal to 1e-3 relative."""
import math

import numpy as np
import pytest

from conftest import LOGIT_TOL, golden_case
from yalm_b200 import api
from yalm_b200 import yalm_file as yf

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("M,N,K", [(128, 128, 64), (1, 8, 8), (200, 320, 512), (512, 1024, 4096), (130, 64, 256), (384, 4096, 1024)])
def test_gemm_f16_tcgen05_matches_numpy(M, N, K):
    rng = np.random.default_rng(M * 7 + N)
    a = (rng.standard_normal((M, K)) / np.sqrt(K)).astype(np.float16)
    w = rng.standard_normal((N, K)).astype(np.float16)
    want = a.astype(np.float32) @ w.astype(np.float32).T
    got = api.gemm_cuda(a, w)
    np.testing.assert_allclose(got, want, atol=2e-3, rtol=1e-3)
    c0 = rng.standard_normal((M, N)).astype(np.float32)
    np.testing.assert_allclose(api.gemm_cuda(a, w, c=c0), want + c0, atol=2e-3, rtol=1e-3)     # residual tail: c += a w^T


@pytest.mark.parametrize("act", [api.ACT_SILU, api.ACT_GELU])
@pytest.mark.parametrize("M,N,K", [(128, 128, 256), (77, 520, 512), (256, 1792, 1024)])
def test_gemm_glu_tcgen05_matches_numpy(M, N, K, act):
    rng = np.random.default_rng(N + K)
    a = (rng.standard_normal((M, K)) / np.sqrt(K)).astype(np.float16)
    w1, w3 = rng.standard_normal((N, K)).astype(np.float16), rng.standard_normal((N, K)).astype(np.float16)
    g = a.astype(np.float32) @ w1.astype(np.float32).T
    u = a.astype(np.float32) @ w3.astype(np.float32).T
    if act == api.ACT_SILU:
        want = g / (1.0 + np.exp(-g)) * u
    else:
        want = 0.5 * g * (1.0 + np.tanh(0.797885 * (g + 0.044715 * g ** 3))) * u
    got = api.gemm_cuda(a, w1, w3=w3, act=act).astype(np.float32)
    np.testing.assert_allclose(got, want, atol=3e-3, rtol=2e-3)


def _oracle_walk(oracle, img, ids, context=0):
    """The token loop: per-position probability of the next token, the last logits."""
    om = oracle.OracleModel(img, context)
    probs = []
    for pos in range(len(ids) - 1):
        lg = om.forward(ids[pos], pos, 1)
        probs.append(oracle.sample_prob(lg, ids[pos + 1]))
    last = om.forward(ids[-1], len(ids) - 1, 1).copy()
    return om, np.array(probs), last


@pytest.mark.parametrize("name,chunks", [("tiny-fp16", [41]), ("tiny-fp16", [17, 24]), ("tiny-kv8-fp16", [30, 11]), ("tiny-gelu-clip-fp16", [41])])
def test_forward_batch_matches_token_loop(oracle, name, chunks):
    fx, img = golden_case(name)
    n = sum(chunks)
    ids = [1] + [(i * 37 + 11) % 500 + 3 for i in range(n - 1)]
    om, want_p, want_last = _oracle_walk(oracle, img, ids)
    m = api.Model(img).cuda()
    assert m.batch_capacity() >= max(chunks)
    s = api.InferenceState(m)
    got_p, pos = [], 0
    for c in chunks:
        tg = ids[pos + 1: pos + c + 1] + [0] * (pos + c + 1 - len(ids))        # the very last position has no target
        got_p += s.forward_batch(ids[pos: pos + c], pos, tg).tolist()
        pos += c
    got_p = np.array(got_p[:-1])
    np.testing.assert_allclose(got_p, want_p, rtol=2e-2, atol=1e-6)
    ok = want_p > 1e-30            # a random-init model can put a target below fp32's range: no log-probability there
    if ok.sum() >= 8:
        ppl = lambda p: math.exp(-np.log(p[ok].astype(np.float64)).mean())
        assert abs(ppl(got_p) - ppl(want_p)) / ppl(want_p) < 1e-3
    assert np.abs(s.logits() - want_last).max() < LOGIT_TOL
    # the KV rows the chunk appended == the rows the decode path appends (fp16; the GEMM sums in another order: <= 2 ulp)
    L, H, D, T = m.n_layers, m.n_kv_heads, m.head_dim, m.max_seq_len
    for layer in range(L):
        for which in (4, 5):
            pre = s.read(which, n * H * D, np.uint16, layer).view(np.float16).astype(np.float32)
            ref = (om.kc if which == 4 else om.vc)[layer][:n].reshape(-1).view(np.float16).astype(np.float32)
            assert np.abs(pre - ref).max() <= 4e-3 * max(1.0, np.abs(ref).max())
    # decode continues from the prefilled cache: same greedy tokens as the oracle continuing its loop
    want_t = []
    lg = want_last
    for i in range(12):
        t = int(np.argmax(lg))
        want_t.append(t)
        lg = om.forward(t, n + i, 1)
    first = int(np.argmax(s.logits()))
    assert first == want_t[0]
    assert s.generate_greedy(first, n, 11) == want_t[1:]
    s.close()
    m.close()


def test_forward_batch_head_dim_128_long_prompt(oracle):
    """Mistral attention geometry, 600-token prompt = one full 512 chunk + a ragged one; the later chunk attends over the
    rows the first one appended."""
    arch = yf.Arch(1024, 2048, 128, 2, 8, 2, 512, 4096, 1000000.0, name="prefill-hd128")
    img = bytes(yf.random_model(arch, "fp16", seed=6))
    ids = [1] + [(i * 131 + 7) % 509 + 3 for i in range(599)]
    om, want_p, want_last = _oracle_walk(oracle, img, ids)
    m = api.Model(img).cuda()
    s = api.InferenceState(m)
    cap = m.batch_capacity()
    got_p, pos = [], 0
    while pos < len(ids):
        c = min(cap, len(ids) - pos)
        tg = ids[pos + 1: pos + c + 1] + [0] * (pos + c + 1 - len(ids))
        got_p += s.forward_batch(ids[pos: pos + c], pos, tg).tolist()
        pos += c
    got_p = np.array(got_p[:-1])
    ppl = lambda p: math.exp(-np.log(p.astype(np.float64)).mean())
    assert abs(ppl(got_p) - ppl(want_p)) / ppl(want_p) < 1e-3
    np.testing.assert_allclose(got_p, want_p, rtol=3e-2, atol=1e-6)
    assert np.abs(s.logits() - want_last).max() < LOGIT_TOL
    s.close()
    m.close()


def test_forward_batch_refuses_what_it_cannot_do():
    _, moe = golden_case("tiny-moe-fp16")
    m = api.Model(moe).cuda()
    assert m.batch_capacity() == 0                 # MoE: the caller walks the tokens
    s = api.InferenceState(m)
    with pytest.raises(api.YalmError):
        s.forward_batch([1, 2, 3], 0)
    s.close()
    m.close()
    fx, img = golden_case("tiny-fp16")
    m = api.Model(img).cuda()
    s = api.InferenceState(m)
    with pytest.raises(api.YalmError):             # would wrap the 64-row ring
        s.forward_batch(list(range(3, 43)), 30)
    s.close()
    m.close()
