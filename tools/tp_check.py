#!/usr/bin/env python3
"""Tensor-parallel parity on N GPUs (run under torchrun): the golden tiny models sharded N ways must reproduce the
reference's greedy tokens and logits (tests/golden/*.npz) on every rank.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29513 tools/tp_check.py
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def main():
    import torch
    import torch.distributed as dist
    from conftest import LOGIT_TOL, golden_case
    from yalm_b200 import api, yalm_file as yf
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    idt = torch.zeros(128, dtype=torch.uint8, device="cuda")
    if rank == 0:
        buf = np.zeros(128, np.uint8)
        api._check(api.lib().yalm_tp_unique_id(buf.ctypes.data), "tp_unique_id")
        idt.copy_(torch.from_numpy(buf))
    dist.broadcast(idt, 0)
    buf = idt.cpu().numpy()
    api._check(api.lib().yalm_tp_init(buf.ctypes.data, rank, world), "tp_init")
    bad = 0
    for name in ["tiny-fp32", "tiny-fp16", "tiny-moe-fp16", "tiny-gelu-clip-fp16"]:
        fx, img = golden_case(name)
        md = yf.load(img)[0]
        if int(md["n_kv_heads"]) % world or int(md["hidden_dim"]) % world or int(md["vocab_size"]) % world:
            if rank == 0:
                print(f"{name}: {md['n_kv_heads']} KV heads / hidden {md['hidden_dim']} / vocab {md['vocab_size']} do not split {world} ways -> SKIP", flush=True)
            continue
        m = api.Model(img, 0).cuda(local, rank, world)
        s = api.InferenceState(m)
        toks, lg = s.generate(fx["prompt"].tolist(), len(fx["tokens"]), keep_logits=True)
        err = float(np.abs(lg[fx["keep"]] - fx["logits"]).max())
        ok = err < LOGIT_TOL and toks == fx["tokens"].tolist()
        print(f"rank {rank}/{world} mega={int(m.uses_megakernel())} {name}: max |dlogit| {err:.2e} tokens {'identical' if toks == fx['tokens'].tolist() else 'DIFFER'} -> {'PASS' if ok else 'FAIL'}", flush=True)
        bad += 0 if ok else 1
        s.close()
        m.close()
    dist.barrier()
    api.lib().yalm_tp_finalize()
    dist.destroy_process_group()
    sys.exit(1 if bad else 0)


if __name__ == "__main__":
    main()
