#!/usr/bin/env python
"""Static evidence for the built library (no GPU needed): per-kernel registers / stack / static shared memory from
`cuobjdump -res-usage`, and counts of the SASS mnemonics that show which hardware paths a kernel uses (tensor-core MMA
on tensor memory, TMA loads/stores, multimem, 128/256-bit global accesses). Mnemonic names follow
/opt/skills/guides/B200_PROFILING.md.

    python tools/static_report.py > profiles/r2_static_sass.txt
"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SO = os.path.join(ROOT, "gke-ray-train_b200", "b200_train", "libb2train.so")

# mnemonic pattern (anchored at the start of the opcode) -> column title
WATCH = [
    ("UTCHMMA", "tcgen05.mma"),
    ("UTCBAR", "tcgen05.commit"),
    ("LDTM", "tcgen05.ld"),
    ("UTMALDG", "TMA load"),
    ("UTMASTG", "TMA store"),
    ("SYNCS", "mbarrier"),
    ("LDGMC", "multimem.ld_reduce"),
    ("STGMC", "multimem.st"),
    ("REDGMC", "multimem.red"),
    (r"LDG\.[A-Z0-9.]*128", "ldg128"),
    (r"LDG\.[A-Z0-9.]*256", "ldg256"),
    (r"STG\.[A-Z0-9.]*128", "stg128"),
    (r"STG\.[A-Z0-9.]*256", "stg256"),
    ("LDL", "local ld"),
    ("STL", "local st"),
]
WATCH = [(re.compile(p), t) for p, t in WATCH]


def run(*cmd):
    return subprocess.run(cmd, check=True, capture_output=True, text=True).stdout


def demangle(names):
    out = run("c++filt", *names).splitlines()
    short = []
    for s in out:
        s = re.sub(r"\(anonymous namespace\)::", "", s)
        s = re.sub(r"^void ", "", s)
        s = re.sub(r"\(.*$", "", s)            # drop the argument list, keep template arguments
        short.append(s)
    return short


def main():
    if not os.path.exists(SO):
        sys.exit(f"{SO} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` first")
    res = {}
    cur = None
    for line in run("cuobjdump", "-res-usage", SO).splitlines():
        m = re.match(r"\s*Function (\S+):", line)
        if m:
            cur = m.group(1)
            continue
        m = re.search(r"REG:(\d+) STACK:(\d+) SHARED:(\d+) LOCAL:(\d+)", line)
        if m and cur:
            res[cur] = tuple(int(x) for x in m.groups())
            cur = None
    counts = collections.defaultdict(collections.Counter)
    cur = None
    for line in run("cuobjdump", "-sass", SO).splitlines():
        m = re.match(r"\s*Function : (\S+)", line)
        if m:
            cur = m.group(1)
            continue
        if cur is None:
            continue
        m = re.match(r"\s*/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z][A-Z0-9_.]*)", line)
        if not m:
            continue
        op = m.group(1)
        counts[cur]["_total"] += 1
        for prefix, _ in WATCH:
            if prefix.match(op):
                counts[cur][prefix] += 1
    names = sorted(res)
    short = demangle(names)
    arch = re.findall(r"arch = (sm_\w+)", run("cuobjdump", "-lelf", SO) + run("cuobjdump", "-elf", SO)[:20000])
    print(f"# {os.path.relpath(SO, ROOT)}: {len(names)} kernels; ELF targets {sorted(set(arch)) or 'n/a'}")
    print("# regs / stack B / static smem B / local B, SASS instruction count, then the watched mnemonics (count > 0 only)")
    for n, s in sorted(zip(names, short), key=lambda t: t[1]):
        r = res[n]
        c = counts[n]
        marks = ", ".join(f"{title}={c[p]}" for p, title in WATCH if c[p])
        print(f"{s:<70s} reg={r[0]:<3d} stack={r[1]:<3d} smem={r[2]:<5d} local={r[3]:<3d} sass={c['_total']:<5d} {marks}")


if __name__ == "__main__":
    main()
