#!/usr/bin/env python
"""DRAM bytes per C-ABI call of one kernel family, from an `ncu --metrics dram__bytes_read.sum,
dram__bytes_write.sum --csv` capture of bench.py (the last `kernels_per_step` matching launches = one
step).  Writes the JSON fragment bench.py reads as roofline.traffic.
usage: ncu_traffic.py dram.csv <family, e.g. cnb_bn_bwd> <kernels_per_step> <calls_per_step> <out.json>"""
import csv, json, sys

def main(path, family, kps, cps, out):
    kps, cps = int(kps), int(cps)
    lines = [l for l in open(path, newline="") if l.startswith('"')]
    per = {}
    for r in csv.DictReader(lines):
        v = float(r["Metric Value"].replace(",", ""))
        u = r["Metric Unit"].lower()
        if r["Metric Name"].startswith("dram__bytes"):
            v *= {"byte": 1, "kbyte": 1e3, "mbyte": 1e6, "gbyte": 1e9}[u]
        per.setdefault(int(r["ID"]), {})[r["Metric Name"]] = v
    if len(per) < kps:
        sys.exit(f"ncu_traffic: {path} holds {len(per)} launches with dram__bytes metrics, expected >= "
                 f"{kps} (empty or truncated capture — re-run ncu; refusing to write {out})")
    ids = sorted(per)[-kps:]
    rd = sum(per[i].get("dram__bytes_read.sum", 0.0) for i in ids)
    wr = sum(per[i].get("dram__bytes_write.sum", 0.0) for i in ids)
    d = {}
    try:
        d = json.load(open(out))
    except Exception:
        pass
    d[family] = {"dram_bytes_per_call": (rd + wr) / cps, "dram_read_bytes_per_step": rd,
                 "dram_write_bytes_per_step": wr, "kernels_per_step": kps, "calls_per_step": cps,
                 "source": "ncu dram__bytes_read.sum + dram__bytes_write.sum, " + path.split("/")[-1]}
    json.dump(d, open(out, "w"), indent=1)
    print(json.dumps(d[family]))

if __name__ == "__main__":
    main(*sys.argv[1:6])
