#!/usr/bin/env python
"""Per-layer-signature table of a bench.py --trace-out file (measurement tooling)."""
import json, sys
from collections import defaultdict

def main(path, top=40):
    t = json.load(open(path))
    agg = defaultdict(lambda: [0, 0.0, 0.0, 0.0])
    for n, a, ms in t["calls"]:
        d = a[0] if a and isinstance(a[0], dict) else {}
        fl = by = 0.0
        if n.startswith("cnb_conv2d"):
            P = (d["H"] + 2 * d["pad_h"] - d["dil_h"] * (d["R"] - 1) - 1) // d["stride_h"] + 1
            Q = (d["W"] + 2 * d["pad_w"] - d["dil_w"] * (d["S"] - 1) - 1) // d["stride_w"] + 1
            fl = 2.0 * d["N"] * P * Q * d["K"] * (d["C"] // d["groups"]) * d["R"] * d["S"]
            by = (d["N"] * d["H"] * d["W"] * d["C"] + d["N"] * P * Q * d["K"]) * 2 + d["K"] * d["C"] // d["groups"] * d["R"] * d["S"] * 2
            key = (n[4:], "C%d K%d k%d s%d H%d g%d" % (d["C"], d["K"], d["R"], d["stride_h"], d["H"], d["groups"]))
        elif n.startswith("cnb_bn"):
            nt = {"cnb_bn_fwd_train": 2, "cnb_bn_fwd_eval": 2, "cnb_bn_bwd": 3}[n] + (1 if d["flags"] & 2 else 0) + \
                (1 if (n == "cnb_bn_bwd" and d["flags"] & 1) else 0)
            by = nt * d["M"] * d["C"] * 2
            key = (n[4:], "C%d M%d f%d" % (d["C"], d["M"], d["flags"]))
        else:
            key = (n[4:], "")
        v = agg[key]
        v[0] += 1; v[1] += ms; v[2] += fl; v[3] += by
    tot = sum(v[1] for v in agg.values())
    print("total traced ms %.3f" % tot)
    for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1])[:top]:
        print("%-14s %-28s x%-2d %7.3f ms %6.3f/call %7.0f TF/s %6.0f GB/s" % (
            k[0], k[1], v[0], v[1], v[1] / v[0], v[2] / (v[1] * 1e-3) / 1e12, v[3] / (v[1] * 1e-3) / 1e9))

if __name__ == "__main__":
    main(sys.argv[1], int(sys.argv[2]) if len(sys.argv) > 2 else 40)
