#!/usr/bin/env python
"""Share-of-step table from an ncu launch list (`ncu --metrics gpu__time_duration.sum --csv --log-file X.csv ...`, optionally
gzipped): per kernel name the number of launches, the summed device time and its share of the captured window.

    python tools/launch_shares.py profiles/r2_bench_launches.csv.gz [--top 30] [--ours]
"""
import argparse
import collections
import csv
import gzip
import io

OURS = ("umma_kernel", "adamw", "sumsq", "clip_coef", "ce_", "splitk", "peer_", "scale_kernel", "gather_windows", "colsum", "land_copy")


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("path")
    ap.add_argument("--top", type=int, default=30)
    ap.add_argument("--ours", action="store_true", help="only this repository's kernels")
    a = ap.parse_args()
    txt = (gzip.open(a.path, "rt") if a.path.endswith(".gz") else open(a.path)).read()
    rows = list(csv.reader(io.StringIO("\n".join(l for l in txt.splitlines() if l.startswith('"')))))
    idx = {h: i for i, h in enumerate(rows[0])}
    agg, total = collections.OrderedDict(), 0.0
    for r in rows[1:]:
        if r[idx["Metric Name"]] != "gpu__time_duration.sum":
            continue
        v = float(r[idx["Metric Value"]].replace(",", "")) * {"ns": 1e-3, "us": 1.0, "ms": 1e3}.get(r[idx["Metric Unit"]], 1.0)
        e = agg.setdefault(r[idx["Kernel Name"]], [0, 0.0])
        e[0] += 1
        e[1] += v
        total += v
    print(f"{sum(c for c, _ in agg.values())} launches, {total / 1e3:.2f} ms of kernel time in the window")
    print("|      us | launches | share | kernel |\n|---:|---:|---:|---|")
    shown = ours = 0.0
    for name, (c, v) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        mine = any(k in name for k in OURS) and "native::" not in name
        ours += v if mine else 0.0
        if (a.ours and not mine) or shown >= a.top:
            continue
        shown += 1
        print(f"| {v:7.0f} | {c:5d} | {100 * v / total:4.1f}% | `{name[:90]}` |")
    print(f"this repository's kernels: {ours:.0f} us = {100 * ours / total:.1f}% of the window")


if __name__ == "__main__":
    main()
