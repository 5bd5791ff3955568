"""World-size-2 check of the tensor-parallel host logic on CPU (gloo): the loader's shard arithmetic
(yalm_tp_shard_plan = yalm::tp_slice, the function Block::cuda / Model::cuda upload with) plus the collective
pattern of SURVEY.md 8(e) — column-parallel wq/wk/wv/w1/w3, row-parallel wo/w2 followed by an all-reduce(sum) of
fp32[dim], vocabulary-split LM head followed by an all-gather — must reproduce the unsharded layer.

No device code runs here: the per-rank arithmetic is numpy on the slices the plan selects; torch.distributed (gloo)
carries the same exchanges NCCL carries on the GPUs.
"""
import ctypes as C
import os
import socket

import numpy as np
import pytest

DIMS = dict(dim=64, hidden=160, head_dim=16, n_heads=8, n_kv_heads=4, vocab=96)
TENSORS = dict(wq=0, wk=1, wv=2, wo=3, w1=4, w2=5, w3=6, wcls=7)


def _plan(lib, tensor, rank, size, bytes_per_weight=4):
    d = DIMS
    dims = (C.c_int * 7)(d["dim"], d["hidden"], d["head_dim"], d["n_heads"], d["n_kv_heads"], d["vocab"], bytes_per_weight)
    out = (C.c_longlong * 4)()
    lib.yalm_tp_shard_plan.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p]
    rc = lib.yalm_tp_shard_plan(dims, TENSORS[tensor], rank, size, out)
    assert rc == 0, rc
    return tuple(out)


def _take(w, plan):
    """Apply a (spitch, offset, width, height) byte plan to a row-major fp32 matrix -> flat fp32 array."""
    spitch, offset, width, height = plan
    raw = np.ascontiguousarray(w, dtype=np.float32).view(np.uint8).reshape(-1)
    return np.concatenate([raw[offset + i * spitch: offset + i * spitch + width] for i in range(height)]).view(np.float32)


def _weights(seed=7):
    d = DIMS
    rng = np.random.default_rng(seed)
    q_dim, kv_dim = d["n_heads"] * d["head_dim"], d["n_kv_heads"] * d["head_dim"]
    shapes = dict(wq=(q_dim, d["dim"]), wk=(kv_dim, d["dim"]), wv=(kv_dim, d["dim"]), wo=(d["dim"], q_dim),
                  w1=(d["hidden"], d["dim"]), w2=(d["dim"], d["hidden"]), w3=(d["hidden"], d["dim"]), wcls=(d["vocab"], d["dim"]))
    w = {k: rng.standard_normal(s).astype(np.float32) / np.sqrt(s[1]) for k, s in shapes.items()}
    x = rng.standard_normal(d["dim"]).astype(np.float32)
    att = rng.standard_normal(q_dim).astype(np.float32)  # stands for the (head-local) attention output
    return w, x, att


def _silu(v):
    return v / (1.0 + np.exp(-v))


def _worker(rank, world, port, q):
    try:
        import torch
        import torch.distributed as dist
        from yalm_b200 import api
        os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
        dist.init_process_group("gloo", rank=rank, world_size=world)
        lib = C.CDLL(api.LIB_PATH)
        d = DIMS
        w, x, att = _weights()
        q_dim, kv_dim = d["n_heads"] * d["head_dim"], d["n_kv_heads"] * d["head_dim"]
        lq, lkv, lhid, lv = q_dim // world, kv_dim // world, d["hidden"] // world, d["vocab"] // world
        sl = {k: _take(w[k], _plan(lib, k, rank, world)) for k in w}
        # column-parallel: this rank's output rows
        q_loc = sl["wq"].reshape(lq, d["dim"]) @ x
        k_loc = sl["wk"].reshape(lkv, d["dim"]) @ x
        v_loc = sl["wv"].reshape(lkv, d["dim"]) @ x
        np.testing.assert_allclose(q_loc, (w["wq"] @ x)[rank * lq:(rank + 1) * lq], rtol=1e-5, atol=1e-5)
        np.testing.assert_allclose(k_loc, (w["wk"] @ x)[rank * lkv:(rank + 1) * lkv], rtol=1e-5, atol=1e-5)
        np.testing.assert_allclose(v_loc, (w["wv"] @ x)[rank * lkv:(rank + 1) * lkv], rtol=1e-5, atol=1e-5)
        # row-parallel wo over this rank's heads + residual folded in by rank 0, then all-reduce(sum)  (Block::enqueue)
        part = sl["wo"].reshape(d["dim"], lq) @ att[rank * lq:(rank + 1) * lq]
        if rank == 0:
            part = part + x
        t = torch.from_numpy(part.astype(np.float32))
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        x1 = t.numpy()
        np.testing.assert_allclose(x1, x + w["wo"] @ att, rtol=1e-4, atol=1e-4)
        # FFN: column-parallel w1/w3, local act*up, row-parallel w2, all-reduce(sum)
        hb = _silu(sl["w1"].reshape(lhid, d["dim"]) @ x1) * (sl["w3"].reshape(lhid, d["dim"]) @ x1)
        part = sl["w2"].reshape(d["dim"], lhid) @ hb
        if rank == 0:
            part = part + x1
        t = torch.from_numpy(part.astype(np.float32))
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        x2 = t.numpy()
        full = x1 + w["w2"] @ (_silu(w["w1"] @ x1) * (w["w3"] @ x1))
        np.testing.assert_allclose(x2, full, rtol=1e-4, atol=1e-4)
        # LM head: vocabulary rows split, all-gather in rank order
        loc = torch.from_numpy((sl["wcls"].reshape(lv, d["dim"]) @ x2).astype(np.float32))
        parts = [torch.empty_like(loc) for _ in range(world)]
        dist.all_gather(parts, loc)
        logits = torch.cat(parts).numpy()
        np.testing.assert_allclose(logits, w["wcls"] @ x2, rtol=1e-4, atol=1e-4)
        dist.barrier()
        dist.destroy_process_group()
        q.put((rank, "ok"))
    except Exception as e:  # noqa: BLE001 - reported to the parent
        q.put((rank, f"{type(e).__name__}: {e}"))


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


@pytest.mark.timeout(300)
def test_tp2_shards_and_collectives_reproduce_the_unsharded_layer():
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = dict(q.get(timeout=240) for _ in procs)
    for p in procs:
        p.join(30)
    assert res == {0: "ok", 1: "ok"}, res


def test_shard_plans_tile_every_tensor_exactly_once():
    from yalm_b200 import api
    lib = C.CDLL(api.LIB_PATH)
    w, _, _ = _weights()
    for world in (1, 2, 4):
        for name, full in w.items():
            seen = np.zeros(full.size * 4, np.int32)
            for rank in range(world):
                spitch, offset, width, height = _plan(lib, name, rank, world)
                for i in range(height):
                    seen[offset + i * spitch: offset + i * spitch + width] += 1
            assert (seen == 1).all(), (name, world)
    # a model that does not split is refused
    dims = (C.c_int * 7)(64, 160, 16, 8, 4, 96, 4)
    out = (C.c_longlong * 4)()
    assert lib.yalm_tp_shard_plan(dims, 0, 0, 3, out) != 0
