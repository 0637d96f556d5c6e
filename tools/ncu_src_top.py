#!/usr/bin/env python
"""Top SASS instructions by warp-stall samples from `ncu --page source --csv` (first kernel whose name
contains the filter).  usage: ncu_src_top.py src.csv[.gz] [filter] [N] [occurrence]"""
import csv, gzip, sys

def main(path, filt="", n=25, occ=0):
    op = gzip.open if path.endswith(".gz") else open
    rows = list(csv.reader(op(path, "rt", newline="")))
    kernels, cur = [], None
    for r in rows:
        if r and r[0] == "Kernel Name":
            cur = {"name": r[1], "rows": []}
            kernels.append(cur)
        elif r and r[0] == "Address":
            cur["hdr"] = r
        elif cur is not None and r:
            cur["rows"].append(r)
    ks = [k for k in kernels if filt in k["name"]]
    k = ks[occ]
    h = {c: i for i, c in enumerate(k["hdr"])}
    tot = sum(int(r[h["# Samples"]]) for r in k["rows"])
    print(k["name"][:120], "total samples", tot, "instructions", len(k["rows"]))
    idx = sorted(range(len(k["rows"])), key=lambda i: -int(k["rows"][i][h["# Samples"]]))[:n]
    stall_cols = [c for c in k["hdr"] if c.startswith("stall_") and "Not Issued" not in c]
    for i in sorted(idx):
        r = k["rows"][i]
        s = int(r[h["# Samples"]])
        st = sorted(((int(r[h[c]]), c) for c in stall_cols), reverse=True)[:2]
        prev = k["rows"][i - 1][h["Source"]].strip()[:40] if i else ""
        print(f"{i:5d} {100 * s / tot:5.1f}%  {r[h['Source']].strip()[:70]:70s} {st[0][1]}:{st[0][0]} {st[1][1]}:{st[1][0]}   <- {prev}")

if __name__ == "__main__":
    a = sys.argv
    main(a[1], a[2] if len(a) > 2 else "", int(a[3]) if len(a) > 3 else 25, int(a[4]) if len(a) > 4 else 0)
