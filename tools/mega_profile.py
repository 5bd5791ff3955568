#!/usr/bin/env python
"""Phase profile of the whole-token decode kernel (decode_mega.cuh): where a token's time goes per phase —
grid-barrier wait, input staging, tile loop, and how much of the tile loop the consumers spent starved for
ring stages (= HBM-bound) — from the kernel's own clock64 stamps.  Usage:
    python tools/mega_profile.py [--workload mistral-7b-fp16-4k] [--layers N] [--pos P] [--out file]"""
import argparse
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from yalm_b200 import api, yalm_file as yf  # noqa: E402
import bench  # noqa: E402

PH = ["qkv", "attn", "wo", "ffn(w1/w3)", "w2", "lm_head"]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="mistral-7b-fp16-4k")
    ap.add_argument("--layers", type=int, default=0)
    ap.add_argument("--pos", type=int, default=4000)
    ap.add_argument("--mhz", type=float, default=1965.0)
    ap.add_argument("--out", default="")
    args = ap.parse_args()
    arch_name, dtype, context, desc = bench.WORKLOADS[args.workload]
    arch = yf.ARCHS[arch_name]
    if args.layers:
        arch = yf.slim(arch, args.layers)
    img = yf.random_model_streamed(arch, dtype, seed=0, device="cuda")
    m = api.Model(img, context).cuda()
    del img
    s = api.InferenceState(m)
    assert m.uses_megakernel()
    for i in range(3):
        s.forward(1, args.pos + i)
    p = s.mega_profile(1, args.pos + 3).astype(np.float64)          # (grid, L, 6, 8)
    us = 1.0 / args.mhz
    lines = []
    t_begin = p[:, 0, 0, 0]
    t_end = p[:, -1, 5, 3]
    lines.append(f"{desc}; pos {args.pos}; grid {p.shape[0]} CTAs; {p.shape[1]} layers")
    lines.append(f"kernel span per CTA (first stamp -> last): mean {np.mean(t_end - t_begin) * us:.1f} us")
    lines.append(f"{'phase':12s} {'barrier':>9s} {'stage':>8s} {'tiles':>8s}   (us per layer, mean over CTAs; max over CTAs in brackets; stage = poll + load + norm of the flagged input vector)")
    tot = np.zeros(4)
    for ph in range(6):
        q = p[:, :, ph, :] if ph < 5 else p[:, -1:, ph, :]
        bar = (q[..., 1] - q[..., 0]) * us
        stg = (q[..., 2] - q[..., 1]) * us
        til = (q[..., 3] - q[..., 2]) * us
        stv = np.zeros_like(til)  # slot 4 is the GPU-wide clock now (the starvation counter needs -DYALM_MEGA_TRACE)
        nl = q.shape[1]
        lines.append(f"{PH[ph]:12s} {bar.mean():9.2f} {stg.mean():8.2f} {til.mean():8.2f}   "
                     f"[{bar.max(0).mean():.2f} {stg.max(0).mean():.2f} {til.max(0).mean():.2f}] x{nl}")
        tot += np.array([bar.mean() * nl, stg.mean() * nl, til.mean() * nl, stv.mean() * nl])
    q = p[:, :, 1, :]
    ok = q[..., 5] > 0
    un = np.where(ok, (q[..., 5] - q[..., 2]) * us, np.nan)
    mp = np.where(ok, (q[..., 6] - q[..., 5]) * us, np.nan)
    mg = np.where(ok & (q[..., 7] > 0), (q[..., 7] - q[..., 5]) * us, np.nan)
    cb = np.where(ok, (q[..., 3] - q[..., 6]) * us, np.nan)
    lines.append(f"attn detail: units {np.nanmean(un):.2f} [{np.nanmax(un, 0).mean():.2f}]  merge+publish {np.nanmean(mp):.2f} "
                 f"[{np.nanmax(mp, 0).mean():.2f}] (merge {np.nanmean(mg):.2f} [{np.nanmax(mg, 0).mean():.2f}])  combine(last CTA) {np.nanmean(cb):.2f} [{np.nanmax(cb, 0).mean():.2f}]")
    # flagged hand-over detail (GEMV phases): slots 4 / 5 = tiles done / staged on the GPU-wide clock (ns), 6 / 7 = SM clock when
    # the sampled words were there / when this thread's words were loaded and verified
    if p[:, 1:, 2, 5].max() > 0:
        prev = {0: 4, 2: None, 3: 2, 4: 3}
        pipe = p[:, 1:, 4, 7].max() > 0 and (p[:, 1:, 4, 6] >= p[:, 1:, 4, 7]).all() and (p[:, 1:, 4, 7] >= p[:, 1:, 4, 2]).all()
        for ph in (0, 3, 4, 2):
            q = p[:, 1:, ph, :]
            if ph == 4 and pipe:  # pipelined hb: no staging step; slots 7 / 6 = warp 0 reaches the last slice / has it
                lines.append(f"hand-over into {PH[ph]:10s}: pipelined (hb slices by bulk copy); warp 0 reaches the last slice "
                             f"{((q[..., 7] - q[..., 2]) * us).mean():5.2f} us into the phase and waits {((q[..., 6] - q[..., 7]) * us).mean():5.2f} us for it "
                             f"(max over CTAs {((q[..., 6] - q[..., 7]) * us).max(0).mean():5.2f})")
                continue
            smp = (q[..., 6] - q[..., 1]) * us
            lod = (q[..., 7] - q[..., 6]) * us
            rst = (q[..., 2] - q[..., 7]) * us
            txt_ = f"hand-over into {PH[ph]:10s}: entry -> polling starts {smp.mean():5.2f}  poll+load+verify {lod.mean():5.2f}  norm/store {rst.mean():5.2f}"
            if prev[ph] is not None:
                e = p[:, 1:, prev[ph], 4] if ph != 0 else p[:, :-1, prev[ph], 4]          # producer phase: tiles done (ns)
                st = q[..., 5]
                last = e.max(0)
                txt_ += (f" | producer end skew (last - mean) {(last - e.mean(0)).mean() / 1e3:5.2f} us, last producer end -> staged: "
                         f"mean {(st - last).mean() / 1e3:5.2f} us, min {(st - last).min(0).mean() / 1e3:5.2f}, max {(st - last).max(0).mean() / 1e3:5.2f}")
            lines.append(txt_)
    lines.append(f"{'token total':12s} {tot[0]:9.1f} {tot[1]:8.1f} {tot[2]:8.1f}   (us per token)")
    txt = "\n".join(lines)
    print(txt)
    if args.out:
        open(args.out, "w").write(txt + "\n")
        np.save(os.path.splitext(args.out)[0] + ".npy", p.astype(np.int64))


if __name__ == "__main__":
    main()
