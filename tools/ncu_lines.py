#!/usr/bin/env python
"""Per-source-line instruction and stall totals from `ncu -i X.ncu-rep --page source --csv --print-source cuda,sass`.
usage: ncu_lines.py file.csv [path-substring] [lo-hi]"""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
want = sys.argv[2] if len(sys.argv) > 2 else ""
lo, hi = (map(int, sys.argv[3].split("-")) if len(sys.argv) > 3 else (0, 10**9))
sec, hdr, data = None, None, {}
for r in rows:
    if len(r) >= 2 and r[0] == "File Path":
        sec = r[1]; data.setdefault(sec, []); continue
    if r and r[0] == "Line No":
        hdr = r; continue
    if sec and hdr and r and r[0].isdigit():
        d = dict(zip(hdr[4:], r[4:]))
        def f(k):
            try:
                return float(d.get(k, "0"))
            except ValueError:
                return 0.0
        data[sec].append((int(r[0]), r[1], f("Instructions Executed"), f("# Samples"), f("stall_long_sb"), f("stall_barrier"), f("stall_short_sb"), f("stall_wait"), f("L2 Theoretical Sectors Global")))
for s, v in data.items():
    print(f"{s}: {sum(x[2] for x in v) / 1e6:.1f} M warp instructions, {sum(x[3] for x in v):.0f} samples")
for s, v in data.items():
    if want and want not in s: continue
    print("==", s)
    for ln, src, inst, smp, lsb, bar, ssb, wt, l2 in v:
        if lo <= ln <= hi and (inst > 0 or smp > 0):
            print(f"{ln:5d} {inst / 1e6:8.2f}M smp {smp:6.0f} lsb {lsb:5.0f} bar {bar:5.0f} ssb {ssb:5.0f} wait {wt:5.0f} l2s {l2 / 1e6:7.2f}M | {src[:110]}")
