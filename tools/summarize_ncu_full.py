#!/usr/bin/env python
"""Condenses `ncu -i X.ncu-rep --page raw --csv` into one line per kernel (mean over captured launches):
duration, DRAM bytes (= roofline `traffic`), registers, occupancy, issue slot and pipe utilisation.

    ncu -i gpurun_out/prof.ncu-rep --page raw --csv > /tmp/raw.csv && python tools/summarize_ncu_full.py /tmp/raw.csv
"""
import csv
import collections
import json
import sys

COLS = [
    ("gpu__time_duration.sum", "us"),
    ("dram__bytes_read.sum", "dram_rd_MB"),
    ("dram__bytes_write.sum", "dram_wr_MB"),
    ("launch__registers_per_thread", "regs"),
    ("launch__grid_size", "grid"),
    ("launch__block_size", "block"),
    ("sm__warps_active.avg.pct_of_peak_sustained_active", "occ%"),
    ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue%"),
    ("sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "alu%"),
    ("sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active", "lsu%"),
    ("sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active", "xu%"),
    ("l1tex__throughput.avg.pct_of_peak_sustained_elapsed", "l1%"),
    ("lts__throughput.avg.pct_of_peak_sustained_elapsed", "l2%"),
    ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram%"),
    ("smsp__inst_executed.sum", "warp_inst"),
]


def to_float(v, unit, want):
    v = float(v.replace(",", "")) if v not in ("", "n/a") else float("nan")
    if want.endswith("_MB"):
        v *= {"byte": 1e-6, "Kbyte": 1e-3, "Mbyte": 1.0, "Gbyte": 1e3}.get(unit, 1.0)
    if want == "us":
        v *= {"ns": 1e-3, "us": 1.0, "ms": 1e3, "s": 1e6}.get(unit, 1.0)
    return v


def main(path):
    rows = list(csv.reader(open(path)))
    hdr, units, data = rows[0], rows[1], rows[2:]
    ix = {h: i for i, h in enumerate(hdr)}
    acc = collections.OrderedDict()
    for r in data:
        name = r[ix["Kernel Name"]].split("(")[0]
        key = (name, r[ix["launch__grid_size"]])
        acc.setdefault(key, []).append(r)
    out = []
    for (name, grid), rs in acc.items():
        d = {"kernel": name, "launches": len(rs)}
        for col, short in COLS:
            if col in ix:
                vals = [to_float(r[ix[col]], units[ix[col]], short) for r in rs]
                d[short] = round(sum(vals) / len(vals), 3)
        out.append(d)
    keys = ["kernel", "launches"] + [s for _, s in COLS]
    print(" ".join(f"{k:>12s}" if k != "kernel" else f"{k:16s}" for k in keys))
    for d in out:
        print(" ".join((f"{d.get(k, ''):>12}" if k != "kernel" else f"{d[k][:16]:16s}") for k in keys))
    if len(sys.argv) > 2:
        json.dump(out, open(sys.argv[2], "w"), indent=1)


if __name__ == "__main__":
    main(sys.argv[1])
