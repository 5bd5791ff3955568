#!/usr/bin/env python3
"""Tensor-parallel smoke run WITHOUT torch.distributed (no NCCL watchdog thread to race the error path): N rank processes,
the NCCL id travels through a file.  Runs a slim model of a bench architecture sharded N ways through the host loop and
the device loop; if the whole-token kernel times out, the rank prints the kernel's failure record (which wait, which CTA).

    python tools/tp_debug.py --world 4 [--arch mistral-7b] [--layers 2] [--timeout-s 15]
"""
import argparse
import os
import subprocess
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def worker(args):
    import torch
    from yalm_b200 import api
    rank, world = args.rank, args.world
    torch.cuda.set_device(rank)
    idf = args.path + ".id"
    buf = np.zeros(128, np.uint8)
    if rank == 0:
        api._check(api.lib().yalm_tp_unique_id(buf.ctypes.data), "tp_unique_id")
        with open(idf + ".tmp", "wb") as f:
            f.write(buf.tobytes())
        os.rename(idf + ".tmp", idf)
    else:
        while not os.path.exists(idf):
            time.sleep(0.05)
        buf = np.frombuffer(open(idf, "rb").read(), np.uint8).copy()
    api._check(api.lib().yalm_tp_init(buf.ctypes.data, rank, world), "tp_init")
    m = api.Model(args.path, 4096).cuda(rank, rank, world)
    s = api.InferenceState(m)
    print(f"rank {rank}: loaded, megakernel={int(m.uses_megakernel())}", flush=True)
    tok, toks = 1, []
    for pos in range(args.steps):
        tok = int(np.argmax(s.forward(tok, pos)))
        toks.append(tok)
    print(f"rank {rank}: host loop ok {toks[:8]}", flush=True)
    dev = s.generate_greedy(tok, args.steps, args.steps)
    print(f"rank {rank}: device loop ok {dev[:8]} ({1000 * s.device_ms / args.steps:.1f} us/token)", flush=True)
    s.close()
    m.close()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--world", type=int, default=2)
    ap.add_argument("--arch", default="mistral-7b")
    ap.add_argument("--layers", type=int, default=2)
    ap.add_argument("--steps", type=int, default=16)
    ap.add_argument("--timeout-s", type=float, default=15.0)
    ap.add_argument("--rank", type=int, default=-1)
    ap.add_argument("--path", default="")
    args = ap.parse_args()
    if args.rank >= 0:
        return worker(args)
    from yalm_b200 import yalm_file as yf
    import torch
    path = f"/dev/shm/yalm_tp_debug_{os.getpid()}.yalm"
    yf.random_model_streamed(yf.slim(yf.ARCHS[args.arch], args.layers), "fp16", seed=0, device="cuda" if torch.cuda.is_available() else "cpu", path=path)
    env = dict(os.environ, YALM_MEGA_TIMEOUT_S=str(args.timeout_s))
    procs = [subprocess.Popen([sys.executable, os.path.abspath(__file__), "--world", str(args.world), "--arch", args.arch, "--layers", str(args.layers),
                               "--steps", str(args.steps), "--rank", str(r), "--path", path], env=env) for r in range(args.world)]
    rcs = []
    for p in procs:
        try:
            rcs.append(p.wait(timeout=180))
        except subprocess.TimeoutExpired:
            p.kill()
            rcs.append(-9)
    for f in (path, path + ".id"):
        if os.path.exists(f):
            os.unlink(f)
    print("exit codes", rcs, flush=True)
    sys.exit(0 if all(rc == 0 for rc in rcs) else 1)


if __name__ == "__main__":
    main()
