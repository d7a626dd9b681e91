#!/usr/bin/env python
"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list: time share per kernel name.
    python tools/launch_summary.py gpurun_out/r1h_launches.csv "header line" > profiles/r1h_launches_summary.txt"""
import collections
import csv
import sys


def main(path, header=""):
    rows = [r for r in csv.reader(l for l in open(path, errors="replace") if l.startswith('"'))]
    head, rows = rows[0], rows[1:]
    ki, vi, ui = head.index("Kernel Name"), head.index("Metric Value"), head.index("Metric Unit")
    agg = collections.defaultdict(lambda: [0.0, 0])
    for r in rows:
        v = float(r[vi].replace(",", ""))
        v *= {"ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3}.get(r[ui], 1e-6)
        a = agg[r[ki][:120]]
        a[0] += v
        a[1] += 1
    total = sum(a[0] for a in agg.values())
    if header:
        print(header)
    print(f"# total {total:.1f} ms over {len(rows)} launches (cold-cache, serialised under ncu: compare SHARES)")
    for name, (ms, n) in sorted(agg.items(), key=lambda kv: -kv[1][0]):
        print(f"{100 * ms / total:6.2f}%  {ms:9.2f} ms  {n:5d}x  {name}")


if __name__ == "__main__":
    main(sys.argv[1], sys.argv[2] if len(sys.argv) > 2 else "")
