"""Per-source-line totals from an ncu report's source page:
    ncu -i REP --page source --csv --print-source cuda,sass > src.csv ; python tools/ncu_lines.py src.csv kernels_fused.cuh [top]
Prints the lines of the named file with the most executed warp instructions, with their stall samples."""
import csv
import sys
from collections import defaultdict


def main():
    path, fname = sys.argv[1], sys.argv[2]
    top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
    cur, hdr = None, None
    inst = defaultdict(float); samp = defaultdict(float); src = {}; long_sb = defaultdict(float); l2 = defaultdict(float)
    total = 0.0
    for row in csv.reader(open(path)):
        if not row:
            continue
        if row[0] == "File Path":
            cur = row[1]; hdr = None; continue
        if row[0] == "Line No":
            hdr = {n: i for i, n in enumerate(row)}; continue
        if hdr is None or cur is None:
            continue
        try:
            ln = int(row[0])
        except ValueError:
            continue
        def f(name):
            try:
                return float(row[hdr[name]] or 0)
            except (KeyError, ValueError, IndexError):
                return 0.0
        v = f("Instructions Executed")
        total += v
        if not cur.endswith(fname):
            continue
        inst[ln] += v; samp[ln] += f("# Samples"); long_sb[ln] += f("stall_long_sb"); l2[ln] += f("L2 Theoretical Sectors Global")
        if row[1]:
            src[ln] = row[1]
    print(f"total warp instructions (all files): {total:.0f};  in {fname}: {sum(inst.values()):.0f}")
    for ln in sorted(inst, key=lambda k: -inst[k])[:top]:
        print(f"{ln:5d} inst {inst[ln]:12.0f} samples {samp[ln]:7.0f} long_sb {long_sb[ln]:7.0f} l2sect {l2[ln]:11.0f} | {src.get(ln, '').strip()[:110]}")


if __name__ == "__main__":
    main()
