"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list: per-kernel totals and (optionally) the
kernel sequence of one transformer layer.   python bench/launch_summary.py gpurun_out/launches.csv [--seq]"""
import collections
import csv
import re
import sys


def load(path):
    rows = list(csv.reader(open(path)))
    hi = [i for i, r in enumerate(rows) if "Kernel Name" in r][0]
    h = rows[hi]
    kn, mv = h.index("Kernel Name"), h.index("Metric Value")
    seq = []
    for r in rows[hi + 1:]:
        try:
            seq.append((r[kn], float(r[mv].replace(",", ""))))
        except (ValueError, IndexError):
            pass
    return seq


def short(n):
    m = re.search(r"(\w+)<([^>]*)>", n)
    return (m.group(1) + "<" + m.group(2) + ">")[:70] if m else n[:70]


def main():
    seq = load(sys.argv[1])
    agg = collections.defaultdict(lambda: [0, 0.0])
    for n, v in seq:
        agg[short(n)][0] += 1
        agg[short(n)][1] += v
    tot = sum(v[1] for v in agg.values())
    print(f"{len(seq)} launches, {tot / 1e6:.3f} ms of kernel time")
    for n, (c, t) in sorted(agg.items(), key=lambda x: -x[1][1])[:30]:
        print(f"{t / 1e3:10.1f} us {c:5d}x {t / c / 1e3:8.1f} us avg {100 * t / tot:5.1f}%  {n}")
    if "--seq" in sys.argv:
        fwd = [i for i, (n, _) in enumerate(seq) if "attn_fwd" in n]
        bwd = [i for i, (n, _) in enumerate(seq) if "attn_bwd" in n]
        if len(fwd) > 3:
            print("--- forward, one layer")
            for n, v in seq[fwd[2] - 2: fwd[3] - 2]:
                print(f"{v / 1e3:8.1f} us  {short(n)}")
        if len(bwd) > 3:
            print("--- backward, one layer")
            for n, v in seq[bwd[2] + 1: bwd[3] + 1]:
                print(f"{v / 1e3:8.1f} us  {short(n)}")


if __name__ == "__main__":
    main()
