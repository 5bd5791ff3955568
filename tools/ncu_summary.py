#!/usr/bin/env python3
"""Condense an .ncu-rep (ncu --set full --import-source on) into a text summary for profiles/.

usage: tools/ncu_summary.py gpurun_out/mega_full.ncu-rep profiles/r1b_mega_full_summary.txt
"""
import csv
import io
import subprocess
import sys

KEYS = [
    "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "dram__bytes_read.sum.per_second",
    "dram__throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_sector_hit_rate.pct", "launch__grid_size",
    "launch__block_size", "launch__registers_per_thread", "launch__shared_mem_per_block_dynamic",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm__warps_active.avg.pct_of_peak_sustained_active",
    "smsp__inst_executed.sum", "sm__icc_request_hit_rate.pct", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
    "smsp__pcsamp_sample_count",
]


def page(rep, name):
    out = subprocess.run(["ncu", "-i", rep, "--page", name, "--csv"], capture_output=True, text=True).stdout
    return list(csv.reader(io.StringIO(out)))


def main():
    rep, dst = sys.argv[1], sys.argv[2]
    raw = page(rep, "raw")
    h, u, row = raw[0], raw[1], raw[2]
    d = dict(zip(h, zip(u, row)))
    lines = [f"# {rep}: kernel {d['Kernel Name'][1]}", "## launch metrics"]
    for k in KEYS:
        if k in d:
            lines.append(f"{k:62s} {d[k][1]:>16s} {d[k][0]}")
    lines.append("## warp stall sampling (all samples)")
    st = [(k, int(float(v[1]))) for k, v in d.items() if k.startswith("smsp__pcsamp_warps_issue_stalled") and not k.endswith("not_issued")]
    tot = sum(v for _, v in st)
    for k, v in sorted(st, key=lambda x: -x[1]):
        lines.append(f"{k.replace('smsp__pcsamp_warps_issue_stalled_', ''):28s} {v:8d} {100.0 * v / max(tot, 1):6.1f}%")
    src = page(rep, "source")
    if len(src) > 2:
        hh = src[1]
        ix = {k: i for i, k in enumerate(hh)}
        rows = []
        for r in src[2:]:
            try:
                rows.append((int(r[ix["# Samples"]] or 0), int(r[ix["Instructions Executed"]] or 0), r[ix["Source"]]))
            except (ValueError, IndexError):
                continue
        lines.append(f"## hottest SASS instructions by samples (of {sum(r[0] for r in rows)} samples, {len(rows)} instructions)")
        for s, e, t in sorted(rows, key=lambda x: -x[0])[:25]:
            lines.append(f"{s:8d} samples {e:12d} exec  {t.strip()[:100]}")
    open(dst, "w").write("\n".join(lines) + "\n")
    print("\n".join(lines[:40]))


if __name__ == "__main__":
    main()
