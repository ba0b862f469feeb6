#!/usr/bin/env python
"""Reads the SASS source page of an ncu report and says where a warp-specialised kernel spends its time.

    ncu -i gpurun_out/x.ncu-rep --page source --csv --print-source sass > /tmp/sass.csv
    python tools/ncu_source_stalls.py /tmp/sass.csv --kernel 'umma_kernel<(int)2' [--instance 0] [--top 30]

For the chosen kernel instance it prints
  * every synchronisation / tensor / TMA instruction with how often it executed and how many warp samples landed on it
    (a `SYNCS...TRYWAIT` pair is a barrier wait: the first try runs once per wait, the second once per failed poll — so
    "first executed N times, loop executed M times" = M polls over the waits that did not succeed at once);
  * the instruction mix of a SASS index range (--range A:B), per execution of a chosen instruction (--per INDEX), e.g.
    instructions per epilogue chunk = range of the chunk body / executions of its LDTM;
  * the most-sampled instructions with their two leading stall reasons.
This is how profiles/README.md "Where F1's time goes" was produced.
"""
import argparse
import collections
import csv
import sys

WATCH = ("SYNCS", "UTCHMMA", "UTCBAR", "LDTM", "UTMALDG", "UTMASTG", "UTMACMDFLUSH", "BAR.SYNC", "UCGABAR", "FENCE",
         "LDGMC", "MEMBAR", "ERRBAR")


def split_kernels(path):
    """ncu concatenates one table per kernel instance: a 'Kernel Name' row, a header row, then instruction rows."""
    rows = list(csv.reader(open(path, newline="")))
    out, i = [], 0
    while i < len(rows):
        if rows[i] and rows[i][0] == "Kernel Name":
            name, hdr, j = rows[i][1], rows[i + 1], i + 2
            while j < len(rows) and not (rows[j] and rows[j][0] == "Kernel Name"):
                j += 1
            out.append((name, hdr, [r for r in rows[i + 2:j] if len(r) == len(hdr)]))
            i = j
        else:
            i += 1
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("csv")
    ap.add_argument("--kernel", default="", help="substring of the kernel name")
    ap.add_argument("--instance", type=int, default=0, help="which matching instance (launch) to read")
    ap.add_argument("--top", type=int, default=25)
    ap.add_argument("--range", default=None, help="A:B SASS index range for the instruction mix")
    ap.add_argument("--per", type=int, default=None, help="normalise the mix by the executions of this SASS index")
    args = ap.parse_args()
    ks = [k for k in split_kernels(args.csv) if args.kernel in k[0]]
    if not ks:
        sys.exit("no kernel matches; names: " + ", ".join(sorted({k[0][:60] for k in split_kernels(args.csv)})))
    name, hdr, data = ks[min(args.instance, len(ks) - 1)]
    isrc, ist, iex = hdr.index("Source"), hdr.index("Warp Stall Sampling (All Samples)"), hdr.index("Instructions Executed")
    stall_cols = [i for i, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
    total = sum(int(r[ist] or 0) for r in data)
    print(f"{name[:100]}\n{len(data)} SASS instructions, {total} warp samples, {len(ks)} matching instance(s)\n")
    print("index  executed  samples  instruction")
    for i, r in enumerate(data):
        s = r[isrc].strip()
        if any(w in s for w in WATCH):
            print(f"{i:5d} {int(r[iex] or 0):9d} {int(r[ist] or 0):8d}  {s[:90]}")
    if args.range:
        a, b = (int(x) for x in args.range.split(":"))
        per = int(data[args.per][iex]) if args.per is not None else 1
        mix, smp, tot = collections.Counter(), collections.Counter(), 0
        for r in data[a:b]:
            toks = r[isrc].strip().split()
            op = toks[1] if toks and toks[0].startswith("@") else (toks[0] if toks else "?")
            e = int(r[iex] or 0)
            mix[op] += e / per
            smp[op] += int(r[ist] or 0)
            tot += e
        print(f"\ninstruction mix of [{a}, {b}) per execution of index {args.per}: {tot / per:.1f} instructions")
        for op, c in mix.most_common(30):
            print(f"  {op:28s} {c:8.1f}   samples {smp[op]}")
    print(f"\ntop {args.top} instructions by samples")
    for i, r in sorted(enumerate(data), key=lambda t: -int(t[1][ist] or 0))[:args.top]:
        why = sorted(((int(r[c] or 0), hdr[c]) for c in stall_cols), reverse=True)[:2]
        print(f"{i:5d} {int(r[ist] or 0):7d} {int(r[iex] or 0):9d}  {r[isrc].strip()[:70]:70s} {why}")


if __name__ == "__main__":
    main()
