#!/usr/bin/env python
"""Digest of `ncu -i X.ncu-rep --page raw --csv`: one block per launch with the metrics the
roofline discussion needs (duration, tensor-pipe activity, DRAM / L2 / shared throughput, issue
slots, top warp-stall reasons).  usage: ncu_raw_digest.py raw.csv [name-filter]"""
import csv, sys

KEYS = [
    ("gpu__time_duration.sum", "duration"),
    ("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed", "tensor pipe active %"),
    ("sm__inst_executed_pipe_tensor_op_utcmma.avg.pct_of_peak_sustained_elapsed", "utcmma inst % of peak"),
    ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "SM throughput %"),
    ("dram__throughput.avg.pct_of_peak_sustained_elapsed", "DRAM throughput %"),
    ("dram__bytes_read.sum", "DRAM read"),
    ("dram__bytes_write.sum", "DRAM write"),
    ("lts__throughput.avg.pct_of_peak_sustained_elapsed", "L2 throughput %"),
    ("lts__t_bytes.sum", "L2 bytes"),
    ("lts__t_sector_hit_rate.pct", "L2 hit %"),
    ("l1tex__data_pipe_lsu_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed", "shared wavefronts %"),
    ("l1tex__throughput.avg.pct_of_peak_sustained_elapsed", "L1/TEX throughput %"),
    ("sm__warps_active.avg.pct_of_peak_sustained_active", "achieved occupancy %"),
    ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue slots busy %"),
    ("launch__grid_size", "grid"), ("launch__block_size", "block"), ("launch__registers_per_thread", "regs"),
    ("launch__shared_mem_per_block_dynamic", "dyn smem"),
]


def main(path, filt=""):
    rows = list(csv.reader(open(path, newline="")))
    hdr = None
    for i, r in enumerate(rows):
        if r and r[0] == "ID":
            hdr, units, body = r, rows[i + 1], rows[i + 2:]
            break
    if hdr is None:
        sys.exit("no header row in " + path)
    col = {n: j for j, n in enumerate(hdr)}
    stall = [n for n in hdr if n.startswith("smsp__average_warps_issue_stalled") and n.endswith("_per_issue_active.ratio")]
    if not stall:
        stall = [n for n in hdr if n.startswith("smsp__average_warp_latency_issue_stalled") or
                 (n.startswith("smsp__warp_issue_stalled") and n.endswith("per_warp_active.pct"))]
    for r in body:
        if len(r) < len(hdr):
            continue
        name = r[col["Kernel Name"]]
        if filt and filt not in name:
            continue
        print(f"== {name[:110]}")
        for k, label in KEYS:
            if k in col:
                print(f"   {label:28s} {r[col[k]]} {units[col[k]]}")
        vals = []
        for n in stall:
            try:
                vals.append((float(r[col[n]].replace(',', '')), n))
            except ValueError:
                pass
        vals.sort(reverse=True)
        for v, n in vals[:4]:
            short = n.replace("smsp__average_warps_issue_stalled_", "").replace("_per_issue_active.ratio", "")
            print(f"   stall {short:32s} {v:.2f}")


if __name__ == "__main__":
    main(sys.argv[1], sys.argv[2] if len(sys.argv) > 2 else "")
