"""Per-kernel summary of an `ncu --metrics gpu__time_duration.sum --csv` launch list of benchmarks/step_launches.py:
takes the last training step (the launches after the last gf::step_begin-delimited boundary) and prints the device
time per kernel name.  Usage: summarize_launches.py launches.csv [out.csv]"""
import collections
import csv
import sys


def load(path):
    rows = list(csv.reader(open(path)))
    hdr = next(i for i, r in enumerate(rows) if r and r[0] == "ID")
    h = rows[hdr]
    ki, vi, ui = h.index("Kernel Name"), h.index("Metric Value"), h.index("Metric Unit")
    out = []
    for r in rows[hdr + 1:]:
        if len(r) > vi:
            v = float(r[vi].replace(",", ""))
            if r[ui] == "ns":
                v /= 1000.0
            elif r[ui] in ("ms", "msecond"):
                v *= 1000.0
            out.append((r[ki], v))
    return out


def last_step(recs):
    # a step starts with the csr->csc histogram (training steps only); take everything from the last one on
    starts = [i for i, (k, _) in enumerate(recs) if "csc_hist" in k]
    if len(starts) >= 2:
        return recs[starts[-1] - 6 if starts[-1] >= 6 else 0:]
    return recs


def main():
    recs = load(sys.argv[1])
    starts = [i for i, (k, _) in enumerate(recs) if "csc_hist" in k]
    step = recs[starts[-2]:starts[-1]] if len(starts) >= 2 else recs
    agg = collections.OrderedDict()
    for k, v in step:
        name = k.split("(")[0].replace("void ", "")[:80]
        agg.setdefault(name, [0, 0.0])
        agg[name][0] += 1
        agg[name][1] += v
    tot = sum(v for _, v in step)
    lines = [("kernel", "launches", "us", "share")]
    for k, (c, v) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        lines.append((k, c, f"{v:.1f}", f"{v / tot:.3f}"))
    lines.append(("TOTAL (serialised, cold caches under ncu)", len(step), f"{tot:.1f}", "1.000"))
    w = csv.writer(open(sys.argv[2], "w", newline="")) if len(sys.argv) > 2 else None
    for ln in lines:
        if w:
            w.writerow(ln)
        print(f"{ln[2]:>10} {ln[1]:>4} {ln[3]:>6}  {ln[0]}")


if __name__ == "__main__":
    main()
