#!/usr/bin/env python
"""Per-kernel counts of the Blackwell-native SASS mnemonics in the shipped library, and the full listing.

    python benchmarks/sass_mnemonics.py [lib.so] -> profiles/sass_mnemonics_r2.csv, profiles/libgfedntm_kernels_r2.sass.gz
"""
import csv
import gzip
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
COLS = [("UTCHMMA", r"\bUTCHMMA"), ("UTCBAR", r"\bUTCBAR"), ("LDTM", r"\bLDTM"), ("UTMALDG", r"\bUTMALDG"),
        ("UTMASTG", r"\bUTMASTG"), ("UBLKCP", r"\bUBLKCP"), ("SYNCS", r"\bSYNCS"), ("MUFU.EX2", r"\bMUFU\.EX2"),
        ("HMMA (legacy)", r"(?<![A-Z])HMMA"), ("ATOM/RED", r"\b(ATOMG?|REDG?|RED)\b|\bATOM\."),
        ("STG.STRONG.SYS", r"\bSTG\.E[.\w]*\.STRONG\.SYS")]


def main():
    lib = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "gfedntm_b200", "ops", "libgfedntm_kernels.so")
    sass = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True, check=True).stdout
    with gzip.open(os.path.join(ROOT, "profiles", "libgfedntm_kernels_r2.sass.gz"), "wt") as f:
        f.write(sass)
    demangle = {}
    names = sorted(set(re.findall(r"Function : (\S+)", sass)))
    out = subprocess.run(["cu++filt"] + names, capture_output=True, text=True).stdout.splitlines()
    for n, d in zip(names, out):
        demangle[n] = d
    rows, cur, counts = [], None, None
    for line in sass.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            if cur is not None:
                rows.append([demangle.get(cur, cur)[:100]] + counts)
            cur, counts = m.group(1), [0] * len(COLS)
            continue
        if cur is None or "/*" not in line:
            continue
        for i, (_, pat) in enumerate(COLS):
            if re.search(pat, line):
                counts[i] += 1
    if cur is not None:
        rows.append([demangle.get(cur, cur)[:100]] + counts)
    with open(os.path.join(ROOT, "profiles", "sass_mnemonics_r2.csv"), "w", newline="") as f:
        w = csv.writer(f)
        w.writerow(["kernel"] + [c for c, _ in COLS])
        w.writerows(rows)
    print(f"{len(rows)} kernels, {sass.count(chr(10))} SASS lines")


if __name__ == "__main__":
    main()
