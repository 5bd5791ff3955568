"""yalm_attn_decode_static: the split-K decode attention with its scratch in a file-scope device array (no allocation in
the entry point).  Collected last on purpose.  Checked through the C ABI against the
existing entry point yalm_attn_decode (caller-owned workspace) and the oracle, on two calls in a row with different
inputs, so contents left in the shared scratch by the first call would show in the second."""
import ctypes as C

import numpy as np
import pytest

from conftest import has_gpu
from yalm_b200 import api

pytestmark = [pytest.mark.gpu, pytest.mark.skipif(not has_gpu(), reason="needs a CUDA device")]

CTL_KVLEN = 4


def f16bits(a):
    return np.asarray(a, np.float32).astype(np.float16).view(np.uint16)


class Dev:
    """Device buffers through the reference's own helpers (upload_cuda / download_cuda_raw / free_cuda)."""

    def __init__(self):
        self.L = api.lib()
        self.L.download_cuda_raw.restype = C.c_void_p
        self.L.download_cuda_raw.argtypes = [C.c_void_p, C.c_size_t]
        self.ptrs = []

    def up(self, a):
        a = np.ascontiguousarray(a)
        p = self.L.upload_cuda(a.ctypes.data, a.nbytes)
        assert p
        self.ptrs.append(p)
        return p

    def down(self, p, n):   # blocking copy on the legacy default stream: ordered after the launches below (stream NULL)
        h = self.L.download_cuda_raw(p, n * 4)
        return np.ctypeslib.as_array(C.cast(h, C.POINTER(C.c_float)), (n,)).copy()

    def free(self):
        for p in self.ptrs:
            self.L.free_cuda(p)
        self.ptrs = []


def case(seed, n_heads, n_kv, hd, max_len, kv_len):
    r = np.random.default_rng(seed)
    k = r.standard_normal((n_kv, max_len, hd)).astype(np.float32)       # head-major ring, as the device keeps it
    v = r.standard_normal((n_kv, max_len, hd)).astype(np.float32)
    q = (0.3 * r.standard_normal(n_heads * hd)).astype(np.float32)
    ctl = np.zeros(8, np.int32)
    ctl[CTL_KVLEN] = kv_len
    return f16bits(k), f16bits(v), q, ctl


def run_both(d, k, v, q, ctl, n_heads, n_kv, hd, max_len):
    L = d.L
    dk, dv, dq, dctl = d.up(k), d.up(v), d.up(q), d.up(ctl)
    n = n_heads * hd
    junk = np.full(n, np.nan, np.float32)
    o_ws, o_st = d.up(junk), d.up(junk)
    ws_bytes = L.yalm_attn_workspace_bytes(n_heads, n_kv, hd)
    ws = d.up(np.zeros(ws_bytes, np.uint8))
    assert L.yalm_attn_decode(o_ws, dk, dv, dq, ws, n_heads, n_kv, hd, max_len, dctl, None) == 0
    assert L.yalm_attn_decode_static(o_st, dk, dv, dq, n_heads, n_kv, hd, max_len, dctl, None) == 0
    return d.down(o_ws, n), d.down(o_st, n)


def test_static_scratch_attention_matches_workspace_entry_twice(oracle):
    d = Dev()
    try:
        assert d.L.yalm_attn_decode_static_max_bytes() == 2 << 20
        # Mistral-7B geometry at a ragged length, then a different shape and length straight after it
        for seed, (n_heads, n_kv, hd, max_len, kv_len) in enumerate([(32, 8, 128, 4096, 3001), (16, 8, 64, 1024, 517),
                                                                     (32, 8, 128, 4096, 7)]):
            k, v, q, ctl = case(seed, n_heads, n_kv, hd, max_len, kv_len)
            want, got = run_both(d, k, v, q, ctl, n_heads, n_kv, hd, max_len)
            assert np.isfinite(got).all()
            np.testing.assert_allclose(got, want, atol=2e-4, rtol=1e-3)     # the tolerance of the flash-decode tests
            assert (got == want).all()                                      # same kernel, same split count: same bits
            # oracle takes the reference's (seq, kv_head, head_dim) layout
            kr = np.ascontiguousarray(k.reshape(n_kv, max_len, hd).transpose(1, 0, 2)).ravel()
            vr = np.ascontiguousarray(v.reshape(n_kv, max_len, hd).transpose(1, 0, 2)).ravel()
            xo, _ = oracle.mha(kr, vr, q, hd, kv_len, max_len, n_heads, n_kv)
            np.testing.assert_allclose(got, xo, atol=2e-4, rtol=1e-3)
    finally:
        d.free()


def test_static_scratch_attention_rejects_oversized_shapes(capfd):
    d = Dev()
    try:
        one = d.up(np.zeros(64, np.float32))
        # 512 KV heads x 8 q-heads each x head_dim 128: the partials alone exceed 2 MiB; 128 KV heads: more than 64 counters
        for n_heads, n_kv in ((4096, 512), (128, 128)):
            rc = d.L.yalm_attn_decode_static(one, one, one, one, n_heads, n_kv, 128, 16, one, None)
            assert rc == -1     # YALM_EINVAL
        assert d.L.yalm_attn_decode_static(None, one, one, one, 32, 8, 128, 16, one, None) == -1
        assert "yalm_attn_decode_static" in capfd.readouterr().err
    finally:
        d.free()
