"""-m "not gpu": the host side of the reference-shaped CLIs — build/main's command line (src/main.cpp:290-394) and its
tokenizer (src/tokenizer.cpp) — and the C++ caller of tests/cpp compiled against csrc/yalm_api.h."""
import os
import subprocess

import numpy as np
import pytest

from conftest import ROOT
from yalm_b200 import api
from yalm_b200 import yalm_file as yf


def _vocab_image(pieces, bos=1, eos=2):
    toks = np.frombuffer(b"\0".join(pieces) + b"\0", np.uint8).copy()
    md = {"bos_token_id": str(bos), "eos_token_id": str(eos)}
    return bytes(yf.build_image({"tokenizer.tokens": toks}, md, {"tokenizer.tokens": "U8"}))


SP_VOCAB = ([b"<unk>", b"<s>", b"</s>"] + [b"<0x%02X>" % i for i in range(256)] +
            [b" ", b"  ", b" t", b"in", b"er", b" the", b" th", b"he", b"e", b"t", b"h", b" there", b"re", b"r",
             "▁".encode(), "▁▁".encode(), b"in", b"<|eot_id|>", b"a", b"ab", b"abc", b"abd"])


@pytest.mark.parametrize("text", [b" there in the", b"abdabcab a", b"xyz\xff\x00q", "▁▁▁ café".encode(), b"", b"in"])
def test_tokenizer_matches_reference_algorithm(oracle, text):
    img = _vocab_image(SP_VOCAB)
    text = text.replace(b"\x00", b"")            # C strings: the CLI cannot carry NUL either
    want = oracle.tokenize(SP_VOCAB, text, 1)
    got = api.tokenize(img, text)
    assert got == want
    assert api.detokenize(img, got) == oracle.detokenize(SP_VOCAB, got, 1)
    assert api.tokenize(img, text, add_bos=False) == want[1:]


def test_tokenizer_duplicate_piece_keeps_last_id(oracle):
    img = _vocab_image(SP_VOCAB)
    assert api.tokenize(img, b"in", add_bos=False) == [SP_VOCAB.index(b"in", SP_VOCAB.index(b"in") + 1)]


def test_tokenizer_synthetic_vocab_round_trip(oracle):
    img = bytes(yf.random_model(yf.ARCHS["tiny"], "fp16", seed=14))
    ids = [1, 17, 300, 42, 5, 3, 511]
    text = b"".join(b" t%d" % i for i in ids[1:])
    assert api.tokenize(img, text) == ids
    assert api.detokenize(img, ids) == b"<s>" + text[1:]      # the space after BOS is stripped (src/tokenizer.cpp:46-49)


@pytest.mark.skipif(not os.path.exists(os.path.join(ROOT, "oracle", "_ref", "libyalm_ref.so")), reason="reference not built")
def test_tokenizer_matches_live_reference(oracle, tmp_path):
    img = _vocab_image(SP_VOCAB)
    # the reference opens a model; give it one with this vocabulary
    tiny = yf.ARCHS["tiny"]
    tensors, hints = {}, {}
    for name, arr, st in yf.random_tensors(tiny, "fp32", 1):
        tensors[name], hints[name] = arr, st
    tensors["tokenizer.tokens"] = np.frombuffer(b"\0".join(SP_VOCAB) + b"\0", np.uint8).copy()
    full = bytes(yf.build_image(tensors, yf.metadata_of(tiny, "fp32"), hints))
    ref = oracle.RefModel(full)
    try:
        for text in [b" there in the", b"abdabcab a", b"xyz\xffq", "▁▁▁ café".encode(), b"in"]:
            want = ref.tokenize(text)
            assert api.tokenize(img, text) == want
            assert api.detokenize(img, want) == ref.detokenize(want)
    finally:
        ref.close()


def _main(*args):
    exe = os.path.join(ROOT, "build", "main")
    if not os.path.exists(exe):
        subprocess.check_call(["make", "-s", "-C", ROOT, "cli"])
    return subprocess.run([exe, *args], capture_output=True, text=True, timeout=60)


def test_main_usage_and_flag_parsing(tmp_path):
    assert _main().returncode == 1
    r = _main("x.yalm", "-i")                      # flag without a value
    assert r.returncode == 1 and "Usage" in r.stderr
    assert _main("x.yalm", "-q", "1").returncode == 1
    assert _main("x.yalm", "-m", "zzz", "-i", "a").returncode == 1
    assert _main("x.yalm", "-m", "completion").returncode == 1          # needs exactly one of -i / -f
    assert _main("x.yalm", "-i", "a", "-f", "b").returncode == 1
    r = _main("x.yalm", "-m", "pass", "-n", "10", "-l", "10")           # passkey position out of range
    assert r.returncode == 1 and "passkey position" in r.stderr
    # -d takes a unique prefix ("c" -> cpu, the first match, as in src/main.cpp:344-357) and cpu is refused: no CPU path
    for d in ("cpu", "c"):
        r = _main("x.yalm", "-d", d, "-i", "a")
        assert r.returncode == 1 and "no CPU fallback" in r.stderr


def test_cpp_caller_compiles_against_the_cpp_api(tmp_path, oracle):
    """tests/cpp/test_kernels.cpp mirrors src/test.cpp:148-206 + the run_completion call sequence against csrc/yalm_api.h."""
    exe = os.path.join(ROOT, "build", "test_kernels")
    subprocess.check_call(["make", "-s", "-C", ROOT, "build/test_kernels"])      # rebuilt whenever a header or the library changed
    assert os.path.exists(exe)
    assert os.path.getmtime(exe) >= os.path.getmtime(os.path.join(ROOT, "yalm_b200", "csrc", "yalm_api.h"))
