#!/usr/bin/env python
"""Per-layer-signature table of a bench.py --trace-out file (measurement tooling).  Uses bench.py's
own algorithmic byte / flop model (bench.call_cost) so the table and the bench line always agree."""
import json, os, sys
from collections import defaultdict

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from bench import call_cost  # noqa: E402


def main(path, top=40):
    t = json.load(open(path))
    agg = defaultdict(lambda: [0, 0.0, 0.0, 0.0])
    for n, a, ms in t["calls"]:
        d = a[0] if a and isinstance(a[0], dict) else {}
        fl, by, _ = call_cost(n, a)
        if n.startswith("cnb_conv2d") and d:
            key = (n[4:], "C%d K%d k%d s%d H%d g%d" % (d["C"], d["K"], d["R"], d["stride_h"], d["H"], d["groups"]))
        elif n.startswith("cnb_bn") and d:
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
