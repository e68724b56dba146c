#!/usr/bin/env python
"""Per-source-line instruction counts from an ncu report (needs -lineinfo and --import-source on):
    ncu -i X.ncu-rep --page source --print-source cuda,sass --csv -k regex:<kernel> > /tmp/src.csv
    python tools/ncu_source_hot.py /tmp/src.csv [top]
"""
import csv
import sys


def main(path, top=30):
    rows = list(csv.reader(open(path)))
    hi = [i for i, r in enumerate(rows) if r and r[0] == "Line No"][0]
    h = rows[hi]
    iline, isrc = 0, 1
    iinst = h.index("Instructions Executed")
    ithr = h.index("Thread Instructions Executed")
    ismp = h.index("# Samples")
    out, tot, tots = [], 0, 0
    for r in rows[hi + 1:]:
        if not r or not r[iline].isdigit():
            continue
        try:
            n, th, s = int(r[iinst]), int(r[ithr]), int(r[ismp])
        except ValueError:
            continue
        tot += n
        tots += s
        out.append((n, th, s, int(r[iline]), r[isrc].strip()[:100]))
    print(f"total warp instructions {tot}, samples {tots}")
    print(f"{'inst%':>6s} {'smp%':>6s} {'thr/inst':>8s} {'line':>5s}  source")
    for n, th, s, ln, src in sorted(out, reverse=True)[:top]:
        print(f"{100 * n / max(tot, 1):6.1f} {100 * s / max(tots, 1):6.1f} {th / max(n, 1):8.1f} {ln:5d}  {src}")


if __name__ == "__main__":
    main(sys.argv[1], int(sys.argv[2]) if len(sys.argv) > 2 else 30)
