"""Summarise per-instruction stall samples from `ncu -i X.ncu-rep --page source --csv --kernel-name regex:K`."""
import csv
import sys


def num(v):
    try:
        return int(float(v))
    except ValueError:
        return 0


rows = list(csv.reader(open(sys.argv[1])))
hdrs = [i for i, r in enumerate(rows) if r and r[0] == "Address"]
for n, hi in enumerate(hdrs[:1]):
    hdr = rows[hi]
    idx = {h: i for i, h in enumerate(hdr)}
    end = hdrs[n + 1] - 1 if n + 1 < len(hdrs) else len(rows)
    data = [r for r in rows[hi + 1:end] if len(r) == len(hdr)]
    tot = sum(num(r[idx["# Samples"]]) for r in data)
    print("kernel", rows[hi - 1][1][:100] if hi else "", "total samples", tot, "instrs", len(data))
    stalls = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
    agg = {s: sum(num(r[idx[s]]) for r in data) for s in stalls}
    for s, v in sorted(agg.items(), key=lambda kv: -kv[1])[:8]:
        print(f"  {s:28s}{v:8d} {100 * v / max(tot, 1):5.1f}%")
    top = sorted(data, key=lambda r: -num(r[idx["# Samples"]]))[: int(sys.argv[2]) if len(sys.argv) > 2 else 20]
    for r in top:
        st = {s: num(r[idx[s]]) for s in stalls}
        m = max(st, key=st.get)
        print(r[idx["# Samples"]].rjust(7), r[idx["Source"]][:88].ljust(88), m, st[m])
