"""-m "not gpu": host-side planning arithmetic of the whole-token kernel (csrc/decode_mega.cu: plan_shapes,
decode_mega.cuh: tile_shape / attn_split) through the C ABI, without a device.  For every BASELINE.json architecture
(single GPU and the tensor-parallel shares) the five GEMV tilings must fit a ring stage, cover every output row exactly
once over the CTAs, and the attention split must hand every (KV head, row chunk) to exactly one CTA with no CTA spanning two
heads."""
import ctypes as C

import numpy as np
import pytest

from yalm_b200 import api
from yalm_b200 import yalm_file as yf

STAGE_BYTES = 8192            # mega.h: kStageBytes
BYTES = {api.F32: 4, api.F16: 2, api.F8E4M3: 1}
EP = {api.F32: 4, api.F16: 8, api.F8E4M3: 16}   # weights per 16-byte load


def tile_plan(dtype, dim, q_dim, kv_dim, hidden, vocab):
    out = (C.c_int * 20)()
    rc = api.lib().yalm_mega_tile_plan(dtype, dim, q_dim, kv_dim, hidden, vocab, out)
    return rc, np.array(out[:]).reshape(5, 4)


def local_dims(arch, tp):
    q = arch.n_heads * arch.head_dim // tp
    kv = arch.n_kv_heads * arch.head_dim // tp
    return arch.dim, q, kv, arch.hidden_dim // tp, arch.vocab_size // tp


CASES = [("mistral-7b", "fp16", 1), ("mistral-7b", "fp8", 1), ("mistral-7b", "fp16", 2), ("mistral-7b", "fp16", 4),
         ("mistral-7b", "fp16", 8), ("llama-3.2-1b", "fp32", 1), ("llama-3-70b", "fp16", 8), ("llama-3-70b", "fp16", 1)]


@pytest.mark.parametrize("name,dt,tp", CASES)
def test_gemv_tilings_fit_a_stage_and_cover_every_row(name, dt, tp):
    arch, dtype = yf.ARCHS[name], api.DTYPE_OF[dt]
    dim, q, kv, hid, voc = local_dims(arch, tp)
    rc, sh = tile_plan(dtype, dim, q, kv, hid, voc)
    assert rc == 0
    B = BYTES[dtype]
    rows = [q + 2 * kv, dim, hid, dim, voc]          # qkv, wo, w1/w3 (pairs of rows), w2, lm head
    K = [dim, q, dim, hid, dim]
    for k in range(5):
        R, KT, nk, G = sh[k]
        assert R in (1, 2, 4) and R * KT * B <= STAGE_BYTES          # one tile = one ring stage
        assert KT * nk == K[k] and KT % EP[dtype] == 0               # K splits tile the input exactly, 16-byte loads
        per_group = R // 2 if k == 2 else R                          # w1/w3 tiles hold R/2 rows of each matrix
        assert G * per_group == rows[k]                              # row groups tile the output exactly
    assert sh[0][0] % 2 == 0                                         # RoPE pairs stay inside a tile
    # row group g of a phase runs on CTA (rotation + g) % grid, the rotation advancing by G per phase: over a layer
    # every CTA's share of the streamed bytes differs by less than one row group per phase
    grid = 148
    load = np.zeros(grid)
    rot = 0
    for k in range(4):
        R, KT, nk, G = sh[k]
        cnt = np.bincount((rot + np.arange(G)) % grid, minlength=grid)
        load += cnt * R * KT * nk * B
        assert cnt.max() - cnt.min() <= 1
        rot = (rot + G) % grid
    assert load.max() / load.mean() < (1.01 if tp == 1 else 1.05)   # TP8 of a 7B model: 3 % (few row groups per CTA)


def test_unsupported_shapes_are_rejected():
    assert tile_plan(api.F16, 4096, 4096, 1024, 14336, 32000)[0] == 0
    assert tile_plan(api.F16, 4096, 4096, 1024, 14337, 32000)[0] != 0       # hidden not a multiple of the load width
    assert tile_plan(7, 4096, 4096, 1024, 14336, 32000)[0] != 0             # unknown weight type
    assert api.lib().yalm_mega_tile_plan(api.F16, 4096, 4096, 1024, 14336, 32000, None) != 0


@pytest.mark.parametrize("ncta,n_kv,nch", [(148, 8, 256), (148, 8, 1), (148, 8, 7), (148, 1, 2048), (148, 2, 5),
                                            (132, 8, 256), (148, 8, 2048), (7, 7, 3), (148, 4, 19)])
def test_attention_split_covers_every_chunk_once(ncta, n_kv, nch):
    owner = -np.ones((n_kv, nch), int)
    out = (C.c_int * 5)()
    heads_of = {}
    for cta in range(ncta):
        assert api.lib().yalm_mega_attn_split(cta, ncta, n_kv, nch, out) == 0
        kvh, c0, nc, k0, nu = out[:]
        assert 0 <= kvh < n_kv and c0 <= cta < c0 + nc          # a CTA belongs to exactly one head's CTA range
        heads_of.setdefault(kvh, set()).add((c0, nc))
        assert (owner[kvh, k0:k0 + nu] == -1).all()
        owner[kvh, k0:k0 + nu] = cta
        assert nu <= -(-nch // nc) + 1                          # balanced: at most one chunk above the mean
    assert (owner >= 0).all()                                   # every (head, chunk) has an owner
    assert all(len(v) == 1 for v in heads_of.values()) and len(heads_of) == n_kv
    # the participants of a head (CTAs that own a chunk) number min(nc, nch): the combine waits for exactly that many
    for kvh, v in heads_of.items():
        (c0, nc), = v
        assert len(set(owner[kvh])) == min(nc, nch)


def test_attention_split_rejects_bad_arguments():
    out = (C.c_int * 5)()
    f = api.lib().yalm_mega_attn_split
    assert f(0, 148, 149, 4, out) != 0      # more KV heads than CTAs: plan() refuses such models too
    assert f(148, 148, 8, 4, out) != 0
    assert f(0, 148, 8, 0, out) != 0
    assert f(0, 148, 8, 4, None) != 0


# ---- tensor-core GEMV geometry and pipelined-hb shares (decode_mega.cuh: mma_block / mma_box_rows / hb_share) -----------------
def mma_plan(rows, unit, cap, grid, cta):
    out = (C.c_int * 4)()
    rc = api.lib().yalm_mega_mma_plan(rows, unit, cap, grid, cta, out)
    return rc, tuple(out[:])


@pytest.mark.parametrize("name,grid", [("mistral-7b", 148), ("mistral-7b", 128), ("llama-3-70b", 148), ("llama-3.2-1b", 148)])
def test_tensor_core_tiles_cover_every_row_once_and_fit_a_stage(name, grid):
    arch = yf.ARCHS[name]
    q, kv = arch.n_heads * arch.head_dim, arch.n_kv_heads * arch.head_dim
    # (rows, unit, cap): pairs of rows in tiles of <= 16; w1 / w3: single rows of each in tiles of <= 8 + 8
    for rows, unit, cap in [(q, 2, 16), (kv, 2, 16), (arch.dim, 2, 16), (arch.hidden_dim, 1, 8), (arch.vocab_size, 2, 16)]:
        nxt, mts, over = 0, set(), 0
        for cta in range(grid):
            rc, (mt, row0, nrows, nt) = mma_plan(rows, unit, cap, grid, cta)
            assert rc == 0
            assert row0 == nxt and nrows % unit == 0          # contiguous blocks, whole units
            assert nt == -(-nrows // mt) and (nt - 1) * mt < nrows <= nt * mt or nrows == 0
            nxt += nrows
            mts.add(mt)
            over += nt * mt - nrows                           # rows fetched twice (they belong to the next CTA)
        assert nxt == rows and len(mts) == 1
        mt = mts.pop()
        assert cap // 2 <= mt <= cap or mt == min(cap, -(-rows // unit // grid) * unit)
        stage = 4 * mt * 128 * (2 if unit == 1 else 1)        # 4 column blocks of 128 bytes per row; w1 / w3: two boxes, second at +4096
        assert stage <= STAGE_BYTES and (unit != 1 or 4 * mt * 128 <= 4096)
        if rows // grid >= cap:                               # blocks of at least one full tile: over-fetch stays small
            assert over <= 0.08 * rows, (rows, mt, over)      # (Mistral-7B on 148 CTAs: ~1 %; wk / wv blocks of 6-8 rows: 16 %)
    assert mma_plan(4096, 2, 16, 148, 148)[0] != 0 and mma_plan(4095, 2, 16, 148, 0)[0] != 0


@pytest.mark.parametrize("name,dt,tp", [("mistral-7b", "fp16", 1), ("llama-3-70b", "fp16", 1), ("mistral-7b", "fp8", 1)])
def test_pipelined_hb_shares_add_up_to_the_slices(name, dt, tp):
    """Every slice of hb (one K split of w2) is complete when the per-CTA shares the w1/w3 epilogues count down add up to the
    slice: for every rotation of the row-group -> CTA map."""
    arch, dtype = yf.ARCHS[name], api.DTYPE_OF[dt]
    rc, sh = tile_plan(dtype, *local_dims(arch, tp))
    assert rc == 0
    rh, groups = max(1, int(sh[2][0]) // 2), int(sh[2][3])     # w1/w3: rows of hb per group, row groups
    kt, nk = int(sh[3][1]), int(sh[3][2])                      # w2: columns per tile = slice, K splits
    assert nk >= 2 and kt * nk == arch.hidden_dim // tp and groups * rh == arch.hidden_dim // tp
    grid = 148
    for rot in (0, 1, 57, 147):
        for s in range(nk):
            total = sum(api.lib().yalm_mega_hb_share(groups, rh, kt, grid, cta, rot, s) for cta in range(grid))
            assert total == kt, (rot, s, total)
    assert api.lib().yalm_mega_hb_share(groups, rh, kt, grid, grid, 0, 0) < 0
