"""Source lines with the most warp-stall samples (and their two dominant stall reasons) from an ncu source page:
    ncu -i REP --page source --csv --print-source cuda,sass > src.csv ; python tools/ncu_stall_lines.py src.csv kernels_fused.cuh [top]"""
import csv
import sys
from collections import defaultdict


def main():
    path, fname = sys.argv[1], sys.argv[2]
    top = int(sys.argv[3]) if len(sys.argv) > 3 else 30
    cur = hdr = None
    samp = defaultdict(float); src = {}; st = defaultdict(lambda: defaultdict(float))
    for row in csv.reader(open(path)):
        if not row:
            continue
        if row[0] == "File Path":
            cur, hdr = row[1], None; continue
        if row[0] == "Line No":
            hdr = {n: i for i, n in enumerate(row)}; continue
        if hdr is None or not cur.endswith(fname):
            continue
        try:
            ln = int(row[0])
        except ValueError:
            continue
        def f(n):
            try:
                return float(row[hdr[n]] or 0)
            except (KeyError, ValueError, IndexError):
                return 0.0
        samp[ln] += f("# Samples")
        for k in hdr:
            if k.startswith("stall_") and "Not Issued" not in k:
                st[ln][k] += f(k)
        if row[1]:
            src[ln] = row[1]
    tot = sum(samp.values())
    print("samples", tot)
    for ln in sorted(samp, key=lambda k: -samp[k])[:top]:
        d = sorted(st[ln].items(), key=lambda kv: -kv[1])[:2]
        print(f"{ln:4d} {samp[ln]:7.0f} {100 * samp[ln] / tot:5.1f}% {d[0][0][6:]}={d[0][1]:.0f} {d[1][0][6:]}={d[1][1]:.0f} | {src.get(ln, '').strip()[:100]}")


if __name__ == "__main__":
    main()
