"""-m gpu: the sm_100a kernels, called through the C ABI (libyalm_b200.so), against the oracle on the same
seeded inputs.  Shapes: the reference's own differential tests (src/test.cpp:148-206), its kernel-bench
shapes (src/test.cpp:307-359), and ragged / edge shapes."""
import numpy as np
import pytest

from yalm_b200 import api
from yalm_b200 import yalm_file as yf

pytestmark = pytest.mark.gpu


def f16bits(a):
    return np.asarray(a, np.float32).astype(np.float16).view(np.uint16)


def rnd(shape, seed, scale=1.0):
    return (np.random.default_rng(seed).standard_normal(shape, dtype=np.float32) * np.float32(scale))


def test_attn_known_answer_gpu():
    """test_attn's vectors (src/test.cpp:84-125) through mha_cuda: head_dim 3 takes the generic kernel."""
    q = np.array([0., 1e4, 0., 0., 0., 1e4], np.float32)
    kv = f16bits([1., 0., 0., 0., 1., 0., 0., 0., 1., -1., 0., 0.])
    xout, att = api.mha_cuda(kv, kv, q, head_dim=3, kv_len=4, max_seq_len=4, n_heads=2, n_kv_heads=1)
    np.testing.assert_allclose(att[0], [0., 1., 0., 0.], atol=1e-4)
    np.testing.assert_allclose(att[1], [0., 0., 1., 0.], atol=1e-4)
    np.testing.assert_allclose(xout, [0., 1., 0., 0., 0., 1.], atol=1e-4)


def test_reference_differential_matmul(oracle):
    """src/test.cpp:158-168: fp32 W (16 x 256), abs eps 1e-4."""
    w, x = rnd((16, 256), 0), rnd(256, 1)
    np.testing.assert_allclose(api.matmul_cuda(x, w, api.F32), oracle.matmul(x, w, None, "F32"), atol=1e-4)


def test_reference_differential_mha(oracle):
    """src/test.cpp:171-188: head_dim 16, 16 heads / 8 KV heads, kv_len = max_seq_len = 4; att and xout."""
    kb, vb, q = f16bits(rnd(4 * 8 * 16, 0)), f16bits(rnd(4 * 8 * 16, 1)), rnd(16 * 16, 2)
    xo, ao = oracle.mha(kb, vb, q, 16, 4, 4, 16, 8)
    xg, ag = api.mha_cuda(kb, vb, q, 16, 4, 4, 16, 8)
    np.testing.assert_allclose(ag, ao, atol=1e-4)
    np.testing.assert_allclose(xg, xo, atol=1e-4)


def test_reference_differential_ffn(oracle):
    """src/test.cpp:191-205: fp32 GELU ffn, dim = hidden = 256, weights scaled 1/sqrt(dim)."""
    x = rnd(256, 0)
    w1, w2, w3 = rnd((256, 256), 1, 1 / 16), rnd((256, 256), 2, 1 / 16), rnd((256, 256), 3, 1 / 16)
    np.testing.assert_allclose(api.ffn_cuda(x, w1, w2, w3, api.ACT_GELU, api.F32),
                               oracle.ffn(x, w1, w2, w3, "gelu", "F32"), atol=1e-4)


@pytest.mark.parametrize("dt", ["F32", "F16", "F8_E4M3"])
@pytest.mark.parametrize("d,n", [(16, 256), (7, 64), (33, 1024), (130, 4096), (64, 14336), (5, 1536)])
def test_gemv_shapes_and_dtypes(oracle, dt, d, n):
    """Ragged row counts (d not a multiple of the rows-per-warp), n below / not a multiple of a 32-lane chunk."""
    w, x = rnd((d, n), d * 7 + n, 1 / np.sqrt(n)), rnd(n, 3)
    if dt == "F32":
        got, want = api.matmul_cuda(x, w, api.F32), oracle.matmul(x, w, None, "F32")
    elif dt == "F16":
        wh = w.astype(np.float16)
        got, want = api.matmul_cuda(x, wh, api.F16), oracle.matmul(x, wh.view(np.uint16), None, "F16")
    else:
        codes, scale = yf.quantize_rows_e4m3(w)
        got, want = api.matmul_cuda(x, codes, api.F8E4M3, scale), oracle.matmul(x, codes, scale, "F8_E4M3")
    np.testing.assert_allclose(got, want, atol=1e-4, rtol=1e-4)


def test_gemv_rejects_unaligned_inner_dim():
    with pytest.raises(api.YalmError):
        api.matmul_cuda(np.zeros(30, np.float32), np.zeros((4, 30), np.float16), api.F16)   # n % 8 != 0


def test_gemv_bench_shapes_fp16(oracle):
    """src/test.cpp:316-332 shapes: 14336 x 4096 and 32000 x 4096 fp16 (checked on a row sample)."""
    for d in (14336, 32000):
        w = rnd((d, 4096), d, 1 / 64).astype(np.float16)
        x = rnd(4096, 1)
        got = api.matmul_cuda(x, w, api.F16)
        rows = np.r_[0:64, d // 2:d // 2 + 64, d - 64:d]
        want = oracle.matmul(x, np.ascontiguousarray(w[rows]).view(np.uint16), None, "F16")
        np.testing.assert_allclose(got[rows], want, atol=1e-4, rtol=1e-4)
        assert np.isfinite(got).all()


@pytest.mark.parametrize("act", ["gelu", "silu"])
@pytest.mark.parametrize("dt", ["F16", "F8_E4M3"])
def test_ffn_fp16_fp8(oracle, act, dt):
    dim, hidden = 512, 1536
    x = rnd(dim, 0)
    w1, w2, w3 = rnd((hidden, dim), 1, dim ** -0.5), rnd((dim, hidden), 2, hidden ** -0.5), rnd((hidden, dim), 3, dim ** -0.5)
    a = api.ACT_GELU if act == "gelu" else api.ACT_SILU
    if dt == "F16":
        h = [w.astype(np.float16) for w in (w1, w2, w3)]
        got = api.ffn_cuda(x, *h, a, api.F16)
        want = oracle.ffn(x, *[w.view(np.uint16) for w in h], act, "F16")
    else:
        q = [yf.quantize_rows_e4m3(w) for w in (w1, w2, w3)]
        got = api.ffn_cuda(x, q[0][0], q[1][0], q[2][0], a, api.F8E4M3, q[0][1], q[1][1], q[2][1])
        want = oracle.ffn(x, q[0][0], q[1][0], q[2][0], act, "F8_E4M3", q[0][1], q[1][1], q[2][1])
    np.testing.assert_allclose(got, want, atol=2e-4, rtol=2e-4)


@pytest.mark.parametrize("head_dim,n_heads,n_kv,kv_len,max_len", [
    (128, 32, 8, 1, 64), (128, 32, 8, 17, 64), (128, 32, 8, 700, 1024), (128, 32, 8, 4096, 4096),
    (64, 32, 8, 513, 600), (128, 64, 8, 300, 300), (32, 8, 2, 64, 64), (128, 8, 8, 100, 128), (128, 4, 1, 2000, 2048),
])
def test_flash_decode_vs_oracle(oracle, head_dim, n_heads, n_kv, kv_len, max_len):
    """Split-K flash-decode (all GQA group sizes, single row to 4k rows, ragged splits) vs attn()."""
    kb = f16bits(rnd(max_len * n_kv * head_dim, 0))
    vb = f16bits(rnd(max_len * n_kv * head_dim, 1))
    q = rnd(n_heads * head_dim, 2, 0.3)
    xo, _ = oracle.mha(kb, vb, q, head_dim, kv_len, max_len, n_heads, n_kv)
    xg, _ = api.mha_cuda(kb, vb, q, head_dim, kv_len, max_len, n_heads, n_kv, want_att=False)
    np.testing.assert_allclose(xg, xo, atol=2e-4, rtol=1e-3)


def test_flash_decode_saturated_softmax(oracle):
    """One dominant key per head (the reference KAT idea at production shape): output = that value row."""
    hd, nh, nkv, L = 128, 32, 8, 1000
    k = rnd((L, nkv, hd), 0, 0.01)
    v = rnd((L, nkv, hd), 1)
    q = np.zeros((nh, hd), np.float32)
    for h in range(nh):
        t = (37 * h + 5) % L
        k[t, h // 4] = 0
        k[t, h // 4, h] = 1.0
        q[h, h] = 1e4
    xg, _ = api.mha_cuda(f16bits(k.ravel()), f16bits(v.ravel()), q.ravel(), hd, L, L, nh, nkv, want_att=False)
    xo, _ = oracle.mha(f16bits(k.ravel()), f16bits(v.ravel()), q.ravel(), hd, L, L, nh, nkv)
    np.testing.assert_allclose(xg, xo, atol=1e-3)
