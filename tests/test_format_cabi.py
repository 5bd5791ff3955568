"""Host-side logic that needs no GPU: the .yalm container, fp8 codec, byte accounting, and that the C-ABI
library loads, exports every symbol include/yalm_b200.h declares, and refuses to compute without a GPU."""
import ctypes
import os
import re

import numpy as np
import pytest

from conftest import ROOT, has_gpu
from yalm_b200 import yalm_file as yf


def test_e4m3_roundtrip_and_rounding():
    t = yf.e4m3_decode_table()
    codes = np.arange(256, dtype=np.uint8)
    ok = ~np.isnan(t) & (codes != 0x80)          # -0 encodes to 0x80 too; skip the sign-of-zero case
    assert (yf.f32_to_e4m3(t[ok]) == codes[ok]).all()
    assert yf.f32_to_e4m3(np.array([1e9, -1e9], np.float32)).tolist() == [0x7E, 0xFE]   # saturate at 448
    # ties to even between 1.0 (0x38) and 1.125 (0x39): 1.0625 -> 0x38
    assert yf.f32_to_e4m3(np.array([1.0625], np.float32)).tolist() == [0x38]


def test_container_roundtrip(tmp_path):
    arch = yf.ARCHS["tiny-moe"]
    for dt in ("fp32", "fp16", "fp8"):
        img = yf.random_model(arch, dt, seed=2)
        p = tmp_path / f"m-{dt}.yalm"
        p.write_bytes(img)
        md, tensors, kinds = yf.load(str(p))
        assert md["dtype"] == dt and int(md["n_experts"]) == 4
        assert tensors["model.layers.1.mlp.w1.weight"].shape == (4, 512, 256)
        if dt == "fp8":
            assert kinds["model.layers.0.attn.wq.weight"] == "F8_E4M3"
            assert tensors["model.layers.0.attn.wq.weight.scale"].shape == (256,)
        md2, t2, _ = yf.load(bytes(img))
        assert md2 == md and all((t2[k] == tensors[k]).all() for k in tensors)


def test_weight_byte_accounting_matches_survey():
    """SURVEY.md section 8(d) / BASELINE.md section 3."""
    assert yf.weight_bytes_per_token(yf.ARCHS["mistral-7b"], "fp16") == 14_221_860_864
    assert yf.weight_bytes_per_token(yf.ARCHS["mixtral-8x7b"], "fp16") == 25_498_247_168
    assert yf.weight_bytes_per_token(yf.ARCHS["llama-3-70b"], "fp16") == 139_008_720_896
    assert yf.weight_bytes_per_token(yf.ARCHS["llama-3.2-1b"], "fp32") == 4_943_265_792
    assert yf.kv_bytes_per_token(yf.ARCHS["mistral-7b"], 4096) == 536_870_912


def test_library_exports_every_declared_symbol():
    from yalm_b200 import api
    hdr = open(os.path.join(ROOT, "include", "yalm_b200.h")).read()
    body = hdr[hdr.index('extern "C"'):]
    declared = set(re.findall(r"\b(yalm_[a-z_0-9]+|upload_cuda|download_cuda_raw|register_cuda_host|free_cuda|"
                              r"unregister_cuda_host|set_cuda_device|init_cuda_stream)\s*\(", body))
    declared -= {"yalm_stream_t"}
    assert declared == set(api.SYMBOLS), declared ^ set(api.SYMBOLS)
    L = ctypes.CDLL(api.LIB_PATH)
    for s in declared:
        assert hasattr(L, s), f"libyalm_b200.so does not export {s}"
    assert "sm_100a" in api.lib().yalm_version().decode()


def test_model_opens_and_validates_without_gpu(tmp_path):
    """Config parsing / tensor validation is host logic; compute must refuse without a GPU (no fallback)."""
    from yalm_b200 import api
    img = bytes(yf.random_model(yf.ARCHS["tiny"], "fp16", seed=1))
    m = api.Model(img)
    assert (m.dim, m.hidden_dim, m.n_layers, m.n_heads, m.n_kv_heads, m.vocab_size, m.max_seq_len) == \
           (256, 512, 2, 8, 2, 512, 64)
    m2 = api.Model(img, context=32)          # -T replaces max_seq_len (reference src/model.cpp:31-36)
    assert m2.max_seq_len == 32
    big = bytes(yf.random_model(yf.slim(yf.ARCHS["mistral-7b"], 1, 512), "fp16", seed=1))
    assert api.Model(big).max_seq_len == 4096      # clamp to 4096 without -T
    if not has_gpu():
        with pytest.raises(api.YalmError) as e:
            m.cuda()
        assert e.value.code == -2
        with pytest.raises(api.YalmError):
            api.matmul_cuda(np.zeros(8, np.float32), np.zeros((2, 8), np.float32))
    with pytest.raises(api.YalmError):
        api.Model(b"\x05\x00\x00\x00\x00\x00\x00\x00{bad}")
