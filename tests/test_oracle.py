"""The oracle (oracle/yalm_oracle.c) pinned against the reference: its only known-answer vector, the golden
fixtures produced by the unmodified reference, and — when oracle/_ref is present — the reference itself."""
import numpy as np
import pytest

from conftest import GOLDEN_CASES, LOGIT_TOL, golden_case


def f16(a):
    return np.asarray(a, np.float32).astype(np.float16).view(np.uint16)


def test_attn_known_answer(oracle):
    """reference src/test.cpp:68-126 (test_attn): 2 q-heads on 1 KV head, head_dim 3, kv_len 4."""
    q = np.array([0., 1e4, 0., 0., 0., 1e4], np.float32)
    kv = f16([1., 0., 0., 0., 1., 0., 0., 0., 1., -1., 0., 0.])
    xout, att = oracle.mha(kv, kv, q, head_dim=3, kv_len=4, max_seq_len=4, n_heads=2, n_kv_heads=1)
    np.testing.assert_allclose(att[0], [0., 1., 0., 0.], atol=1e-4)
    np.testing.assert_allclose(att[1], [0., 0., 1., 0.], atol=1e-4)
    np.testing.assert_allclose(xout, [0., 1., 0., 0., 0., 1.], atol=1e-4)


def test_moe_gate_tie_rule(oracle):
    """src/infer.cpp:100-132: strict '>' => lowest index wins ties; weights = softmax over the selected."""
    w, e = oracle.moe_gate([1.0, 3.0, 3.0, 0.5], 2)
    assert e.tolist() == [1, 2]
    np.testing.assert_allclose(w, [0.5, 0.5], atol=1e-6)
    w, e = oracle.moe_gate([0.0, 2.0, 1.0, -1.0], 2)
    assert e.tolist() == [1, 2]
    np.testing.assert_allclose(w, np.exp([0., -1.]) / np.exp([0., -1.]).sum(), atol=1e-6)


def test_matmul_fp8_is_fp32_path_on_dequantised_weights(oracle):
    from yalm_b200 import yalm_file as yf
    rng = np.random.default_rng(0)
    w = rng.standard_normal((24, 64), dtype=np.float32)
    x = rng.standard_normal(64, dtype=np.float32)
    codes, scale = yf.quantize_rows_e4m3(w)
    deq = yf.e4m3_to_f32(codes) * scale[:, None]
    got = oracle.matmul(x, codes, scale, "F8_E4M3")
    want = oracle.matmul(x, deq, None, "F32")
    np.testing.assert_allclose(got, want, rtol=1e-5, atol=1e-5)
    np.testing.assert_allclose(got, w @ x, rtol=0.1, atol=0.35)


@pytest.mark.parametrize("name", GOLDEN_CASES)
def test_oracle_matches_reference_golden(oracle, name):
    """Greedy tokens identical to the reference's, logits within the north-star tolerance, for 72 steps
    (max_seq_len is 64: the last 13 steps run on the ring buffer with the two attention sinks)."""
    fx, img = golden_case(name)
    m = oracle.OracleModel(img)
    toks, logits = m.generate(fx["prompt"].tolist(), len(fx["tokens"]), keep_logits=True)
    assert toks == fx["tokens"].tolist()
    assert np.abs(logits[fx["keep"]] - fx["logits"]).max() < LOGIT_TOL / 4


def test_oracle_matches_live_reference(oracle):
    """Same comparison against oracle/_ref run live, on a model that is not among the fixtures."""
    if not oracle.ref_available():
        pytest.skip("oracle/_ref not built (needs /root/reference)")
    from yalm_b200 import yalm_file as yf
    img = bytes(yf.random_model(yf.ARCHS["tiny-moe"], "fp16", seed=3))
    o, r = oracle.OracleModel(img), oracle.RefModel(img)
    to, lo = o.generate([1, 5, 9], 40, True)
    tr, lr = r.generate([1, 5, 9], 40, True)
    r.close()
    assert to == tr
    assert np.abs(lo - lr).max() < LOGIT_TOL / 4
