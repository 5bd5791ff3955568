"""-m gpu: the reference-shaped command lines on a real device — `build/test -k <name>` (README.md:79-80 /
src/test.cpp:361-383), `build/main <ckpt> -d cuda ...` (src/main.cpp:290-429) and the C++ caller of tests/cpp — checked
against the reference's golden tokens and the oracle."""
import math
import os
import re
import subprocess

import numpy as np
import pytest

from conftest import ROOT, golden_case
from yalm_b200 import yalm_file as yf

pytestmark = pytest.mark.gpu


def _exe(name):
    p = os.path.join(ROOT, "build", name)
    if not os.path.exists(p):
        subprocess.check_call(["make", "-s", "-C", ROOT, "cli"])
    return p


def _run(*cmd, timeout=300):
    r = subprocess.run(list(cmd), capture_output=True, text=True, timeout=timeout, cwd=ROOT)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    return r.stdout


@pytest.mark.parametrize("name,flag", [("matmul", "-k"), ("matmul-wide", "-k"), ("mha", "-k"), ("ffn", "-k"), ("stream", "-k"),
                                       ("matmul", "-bk")])
def test_kernel_bench_cli(name, flag):
    out = _run(_exe("test"), flag, name)
    m = re.search(rf"^{re.escape(name)}: ([0-9.]+) us/launch, ([0-9.]+) GB/s", out, re.M)
    assert m, out
    assert float(m.group(2)) > 500.0          # bandwidth-bound kernels on a B200: anything sane is far above this


def test_kernel_bench_cli_rejects_unknown():
    r = subprocess.run([_exe("test"), "-k", "nope"], capture_output=True, text=True, timeout=120)
    assert r.returncode != 0


def _ids(text):
    """Synthetic vocabulary (yalm_file.synthetic_tokens): "<unk>", "<s>", "</s>", then " t<id>"."""
    special = {"<unk>": 0, "<s>": 1, "</s>": 2}
    return [special[p] if p in special else int(p.strip()[1:]) for p in re.findall(r"<unk>|<s>|</s>| ?t\d+", text)]


def _until_eos(toks):
    """The completion loop stops after printing an end-of-sequence token (src/main.cpp:107-109)."""
    return toks[:toks.index(2) + 1] if 2 in toks else toks


def _completion_ids(out):
    body = out.split("Encoding stats:", 1)[1].split("\n\n", 1)[1].split("Generation stats:", 1)[0]
    return body


def test_main_completion_greedy_matches_reference_golden(tmp_path):
    fx, img = golden_case("tiny-fp16")
    path = tmp_path / "tiny.yalm"
    path.write_bytes(img)
    prompt = "".join(f" t{t}" for t in fx["prompt"].tolist()[1:])        # BOS (id 1) is added by the tokenizer
    want = fx["tokens"].tolist()
    n = 24
    out = _run(_exe("main"), str(path), "-d", "cuda", "-m", "completion", "-i", prompt, "-n", str(n), "-t", "0")
    assert "Using CUDA" in out and "Generation stats:" in out and "bandwidth:" in out
    assert _ids(_completion_ids(out)) == _until_eos(want[:n])
    # prefix matching of -m / -d and the prompt file flag
    pf = tmp_path / "prompt.txt"
    pf.write_text(prompt)
    out2 = _run(_exe("main"), str(path), "-d", "cu", "-m", "c", "-f", str(pf), "-n", str(n), "-t", "0")
    assert _ids(_completion_ids(out2)) == _until_eos(want[:n])


def test_main_sliding_window_flag(tmp_path, oracle):
    """-T 24 on the golden model: the ring wraps during a 60-token completion (src/model.cpp:31-36, infer.cpp:483-485)."""
    fx, img = golden_case("tiny-fp16")
    path = tmp_path / "tiny.yalm"
    path.write_bytes(img)
    prompt_ids = fx["prompt"].tolist()
    want = oracle.OracleModel(img, context=24).generate(prompt_ids, 60)
    out = _run(_exe("main"), str(path), "-T", "24", "-i", "".join(f" t{t}" for t in prompt_ids[1:]), "-n", "60", "-t", "0")
    assert _ids(_completion_ids(out)) == _until_eos(want)


def test_main_perplexity_matches_oracle(tmp_path, oracle):
    fx, img = golden_case("tiny-fp16")
    path = tmp_path / "tiny.yalm"
    path.write_bytes(img)
    ids = [1] + [(i * 37 + 11) % 500 + 3 for i in range(40)]
    om = oracle.OracleModel(img)
    lps = []
    for pos in range(len(ids) - 1):
        lg = om.forward(ids[pos], pos, 1)
        lps.append(math.log(oracle.sample_prob(lg, ids[pos + 1])))
    want = math.exp(-sum(lps) / len(lps))
    out = _run(_exe("main"), str(path), "-m", "perplexity", "-i", "".join(f" t{t}" for t in ids[1:]))
    m = re.search(r"perplexity: ([0-9.e+]+) ", out)
    assert m, out
    assert abs(float(m.group(1)) - want) / want < 1e-3
    assert f"{len(ids) - 1} tokens" in out


def test_main_temperature_sampling_runs(tmp_path):
    fx, img = golden_case("tiny-fp16")
    path = tmp_path / "tiny.yalm"
    path.write_bytes(img)
    out = _run(_exe("main"), str(path), "-i", " t17 t300", "-n", "12", "-t", "0.8")
    assert "Generation stats:" in out


def test_cpp_caller_runs(tmp_path):
    """tests/cpp/test_kernels.cpp (compiled by the CPU suite): src/test.cpp:148-206 against the oracle + the C++ API loop."""
    exe = os.path.join(ROOT, "build", "test_kernels")
    if not os.path.exists(exe):
        pytest.skip("build/test_kernels not built (run the CPU suite first)")
    fx, img = golden_case("tiny-fp16")
    path = tmp_path / "tiny.yalm"
    path.write_bytes(img)
    out = _run(exe, str(path))
    for what in ("ok matmul", "ok mha att", "ok mha xout", "ok ffn"):
        assert what in out
    toks = [int(x) for x in re.search(r"tokens:((?: \d+)+)", out).group(1).split()]
    assert toks == fx["tokens"].tolist()[:8]
