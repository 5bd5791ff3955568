"""Generates tests/golden/convert_golden.json by running the UNMODIFIED reference converter (/root/reference/convert.py) on
the tiny HuggingFace-layout checkpoints of tests/hf_fixture.py:  python tests/golden/make_convert_golden.py
Per (case, dtype): the metadata dict and, per tensor, safetensors dtype + shape + sha256 of its bytes.  (fp8 is not
recorded: the reference casts to unscaled F8_E5M2, which has no arithmetic anywhere — SURVEY.md Q1.)"""
import hashlib
import json
import os
import subprocess
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import hf_fixture  # noqa: E402
from yalm_b200 import yalm_file as yf  # noqa: E402


def main():
    out = {}
    with tempfile.TemporaryDirectory() as tmp:
        for case in hf_fixture.CASES:
            d = os.path.join(tmp, case)
            hf_fixture.write_hf_dir(d, case)
            for dtype in ("fp16", "fp32"):
                dst = os.path.join(tmp, f"{case}-{dtype}.yalm")
                subprocess.check_call([sys.executable, "/root/reference/convert.py", dst, d, "--dtype", dtype], stdout=subprocess.DEVNULL)
                md, tensors, kinds = yf.load(dst)
                out[f"{case}-{dtype}"] = {"metadata": md, "tensors": {
                    n: {"dtype": kinds[n], "shape": list(a.shape), "sha256": hashlib.sha256(a.tobytes()).hexdigest()} for n, a in tensors.items()}}
    with open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "convert_golden.json"), "w") as f:
        json.dump(out, f, indent=0, sort_keys=True)
    print({k: len(v["tensors"]) for k, v in out.items()})


if __name__ == "__main__":
    main()
