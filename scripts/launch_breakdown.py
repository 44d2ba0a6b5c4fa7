#!/usr/bin/env python
"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list: total kernel time and the top
kernels by share.  Usage: launch_breakdown.py launches.csv [top_n] [first_index last_index]"""
import collections
import csv
import re
import sys


def load(path):
    rows = list(csv.reader(open(path)))
    for i, r in enumerate(rows):
        if "Kernel Name" in r:
            h, start = r, i
            break
    ki, vi, ui = h.index("Kernel Name"), h.index("Metric Value"), h.index("Metric Unit")
    seq = []
    for r in rows[start + 1:]:
        if len(r) <= vi:
            continue
        try:
            v = float(r[vi].replace(",", ""))
        except ValueError:
            continue
        v *= {"ns": 1.0, "us": 1e3, "ms": 1e6, "s": 1e9}.get(r[ui], 1.0)
        seq.append((re.sub(r"\s+", " ", r[ki]), v))
    return seq


def main():
    seq = load(sys.argv[1])
    top = int(sys.argv[2]) if len(sys.argv) > 2 else 25
    if len(sys.argv) > 4:
        seq = seq[int(sys.argv[3]):int(sys.argv[4])]
    agg = collections.defaultdict(lambda: [0, 0.0])
    for n, v in seq:
        k = re.sub(r"<.*", "", n)[:64]
        if "elementwise" in n:
            m = re.search(r"native::(\w*Functor\w*|\w+_kernel_cuda)", n[30:])
            k = "aten elementwise: " + (m.group(1) if m else "?")
        agg[k][0] += 1
        agg[k][1] += v
    tot = sum(v for _, v in agg.values())
    print(f"launches {len(seq)}  total {tot / 1e6:.3f} ms")
    for n, (c, v) in sorted(agg.items(), key=lambda x: -x[1][1])[:top]:
        print(f"{v / 1e3:10.1f} us {100 * v / tot:5.1f}% x{c:4d}  {n}")


if __name__ == "__main__":
    main()
