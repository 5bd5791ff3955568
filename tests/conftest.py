import hashlib
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


def has_gpu() -> bool:
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False


@pytest.fixture(scope="session")
def oracle():
    from oracle import oracle as O
    O.build(ref=False)          # compiles oracle/liboracle.so if stale; building the checker is not using it
    return O


def golden_case(name):
    """-> (fixture dict, model image bytes) ; the image is regenerated from the seed and sha256-checked."""
    from yalm_b200 import yalm_file as yf
    fx = dict(np.load(os.path.join(GOLDEN, name + ".npz")))
    arch, dt = name.rsplit("-", 1)
    img = bytes(yf.random_model(yf.ARCHS[arch], dt, seed=int(fx["seed"])))
    assert hashlib.sha256(img).digest() == fx["sha256"].tobytes(), "random_model() no longer reproduces the golden model"
    return fx, img


GOLDEN_CASES = ["tiny-fp32", "tiny-fp16", "tiny-moe-fp32", "tiny-moe-fp16", "tiny-gelu-clip-fp32", "tiny-gelu-clip-fp16",
                "tiny-kv8-fp16", "tiny-kv8-fp32"]

# North-star tolerance: logits within max |delta| <= 2e-2 of the reference CPU path (BASELINE.json).
LOGIT_TOL = 2e-2
