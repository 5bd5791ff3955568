#!/usr/bin/env python
"""Offline analysis of the whole-token kernel's event trace (CTA 0; library built with -DYALM_MEGA_TRACE, file written by
tools/mega_profile.py under YALM_MEGA_TRACE=<file>): per phase of one layer, how long tiles take to arrive, how long the
owning warp lets a landed tile wait, how long it works on it, and what the warps do between tiles.
    python tools/mega_trace.py gpurun_out/trace.bin [--layer L] [--mhz 1965]"""
import argparse
import sys

import numpy as np

PH = ["qkv", "attn", "wo", "ffn", "w2", "lm"]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("file")
    ap.add_argument("--layer", type=int, default=3)
    ap.add_argument("--mhz", type=float, default=1965.0)
    args = ap.parse_args()
    raw = np.fromfile(args.file, dtype=np.int64)
    n = int(raw[0])
    ev = raw[1:1 + 2 * min(n, (len(raw) - 1) // 2)].reshape(-1, 2)
    tag, clk = ev[:, 0], ev[:, 1]
    typ, warp, idx = (tag >> 56) & 0xff, (tag >> 48) & 0xff, tag & 0xffffffff
    us = 1.0 / args.mhz
    t0 = clk.min()
    # phase stamps: type 5, idx = layer * 64 + phase * 8 + slot
    st = {}
    for i in np.nonzero(typ == 5)[0]:
        st[int(idx[i])] = clk[i]
    print(f"{n} events recorded ({len(ev)} kept)")
    L = args.layer
    issued = {int(idx[i]): clk[i] for i in np.nonzero(typ == 1)[0]}
    wbeg = {int(idx[i]): (clk[i], int(warp[i])) for i in np.nonzero(typ == 2)[0]}
    wend = {int(idx[i]): clk[i] for i in np.nonzero(typ == 3)[0]}
    rel = {int(idx[i]): clk[i] for i in np.nonzero(typ == 4)[0]}
    for ph in range(5):
        k0, k1, k2, k3 = (st.get(L * 64 + ph * 8 + s) for s in range(4))
        if k0 is None or k3 is None:
            continue
        nxt = st.get(L * 64 + (ph + 1) * 8 + 3) if ph < 4 else st.get((L + 1) * 64 + 3)
        # tiles of this phase = ring indices whose consumer wait began between stamp 2 (staged) and stamp 3 of warp 0 ... use all warps:
        tiles = sorted(i for i, (c, w) in wbeg.items() if k1 <= c <= k3 + 4000 and i in wend and i in rel)
        # keep the contiguous block around the median (ring indices of a phase are contiguous)
        if not tiles:
            continue
        lat = np.array([(wend[i] - issued[i]) * us for i in tiles if i in issued])
        waited = np.array([(wend[i] - wbeg[i][0]) * us for i in tiles])
        work = np.array([(rel[i] - wend[i]) * us for i in tiles])
        ahead = np.array([(wbeg[i][0] - issued[i]) * us for i in tiles if i in issued])
        print(f"layer {L} {PH[ph]:5s}: barrier {(k1 - k0) * us:5.2f} stage {(k2 - k1) * us:5.2f} tiles {(k3 - k2) * us:6.2f} us | {len(tiles):4d} tiles: "
              f"issue->landed-or-claimed {np.median(lat):5.2f} (p90 {np.percentile(lat, 90):5.2f}) | consumer waited {waited.mean():5.2f} (median {np.median(waited):5.2f}) | "
              f"work per tile {work.mean():5.3f} (median {np.median(work):5.3f}, p90 {np.percentile(work, 90):5.3f}) | issued before claim {np.median(ahead):5.2f}")
        # per warp: timeline gaps between release of one tile and wait-begin of the next (epilogue + bookkeeping)
        gaps = []
        for w in range(11):
            mine = [i for i in tiles if wbeg[i][1] == w]
            mine.sort(key=lambda i: wbeg[i][0])
            for a, b in zip(mine, mine[1:]):
                gaps.append((wbeg[b][0] - rel[a]) * us)
        if gaps:
            g = np.array(gaps)
            print(f"            between tiles of a warp (release -> next wait): mean {g.mean():5.3f} median {np.median(g):5.3f} p90 {np.percentile(g, 90):5.3f} max {g.max():5.2f}; "
                  f"sum of work over warps {work.sum():6.1f} us, of waits {waited.sum():6.1f}, of gaps {g.sum():6.1f}")


if __name__ == "__main__":
    main()
