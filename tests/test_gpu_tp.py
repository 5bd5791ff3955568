"""-m gpu (needs >= 2 GPUs, skipped otherwise): tensor-parallel parity on real devices.  Spawns tests/tp_worker.py under
torchrun with world 2 / 4 / 8 and both tensor-parallel paths: the whole-token kernel with the in-kernel all-reduce over
NVLink peer memory, and the per-op kernels + one-shot all-reduce kernel (YALM_TP_MEGA=0).  The reference is single-GPU
(src/model.cpp:385-386); parity = the reference's golden tokens / logits on every rank."""
import os
import subprocess
import sys

import pytest

from conftest import ROOT

pytestmark = pytest.mark.gpu


def _gpus():
    import torch
    return torch.cuda.device_count()


@pytest.mark.parametrize("mega", [1, 0])
@pytest.mark.parametrize("world", [2, 4, 8])
def test_tp_parity(world, mega):
    if _gpus() < world:
        pytest.skip(f"needs {world} GPUs, have {_gpus()}")
    env = dict(os.environ, YALM_TP_MEGA=str(mega), OMP_NUM_THREADS="4")
    port = 29600 + world * 2 + mega
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}", "--master-addr",
                        "127.0.0.1", "--master-port", str(port), os.path.join(ROOT, "tests", "tp_worker.py")],
                       capture_output=True, text=True, timeout=600, env=env, cwd=ROOT)
    sys.stdout.write(r.stdout[-4000:])
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert "PASS" in r.stdout and "FAIL" not in r.stdout


def test_expert_parallel_parity():
    """YALM_MOE_EP=1: rank r holds whole experts e = r, r + 2 of the 4-expert golden model (the measured
    alternative); same golden tokens / logits, bit-identical on both ranks."""
    if _gpus() < 2:
        pytest.skip("needs 2 GPUs")
    env = dict(os.environ, YALM_MOE_EP="1", OMP_NUM_THREADS="4")
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
                        "--master-port", "29641", os.path.join(ROOT, "tests", "tp_worker.py")], capture_output=True, text=True, timeout=600,
                       env=env, cwd=ROOT)
    sys.stdout.write(r.stdout[-3000:])
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert "PASS" in r.stdout and "FAIL" not in r.stdout
