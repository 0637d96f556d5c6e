#!/usr/bin/env python
"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list into a per-kernel table
(markdown).  usage: ncu_summary.py launches.csv "<command line>" > profiles/rNN_launches_summary.md"""
import csv, re, sys
from collections import defaultdict

def main(path, cmd=""):
    rows = []
    with open(path, newline="") as f:
        lines = [l for l in f if l.startswith('"')]
    for r in csv.DictReader(lines):
        if r.get("Metric Name") != "gpu__time_duration.sum":
            continue
        v = float(r["Metric Value"].replace(",", ""))
        if r["Metric Unit"] in ("ns", "nsecond"):
            v /= 1e3
        elif r["Metric Unit"] in ("ms", "msecond"):
            v *= 1e3
        name = r["Kernel Name"]
        name = re.sub(r"^void ", "", name)
        name = re.sub(r"\(.*$", "", name)
        rows.append((name, v))
    agg = defaultdict(lambda: [0, 0.0])
    for n, v in rows:
        agg[n][0] += 1
        agg[n][1] += v
    tot = sum(v[1] for v in agg.values())
    print(f"Command: `{cmd}`\n")
    print("| kernel | launches | total us | avg us | share |\n|---|---|---|---|---|")
    for n, (c, t) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print(f"| `{n[:90]}` | {c} | {t:.1f} | {t / c:.1f} | {100 * t / tot:.1f}% |")
    print(f"\nTotal {tot:.1f} us over {len(rows)} launches.")

if __name__ == "__main__":
    main(sys.argv[1], sys.argv[2] if len(sys.argv) > 2 else "")
