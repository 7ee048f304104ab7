#!/usr/bin/env python
"""Static SASS / resource table of fdtd_b200/libfdtd_b200.so: `cuobjdump -sass` per function joined with the register / spill lines of the
last build (fdtd_b200/csrc/build.log, written by `python -m fdtd_b200.build`).  usage: sass_extract.py [kernel-name-substring ...] > table.md"""
import re
import subprocess
import sys

ROOT = __file__.rsplit("/tools/", 1)[0]
MN = ["UTMALDG", "UBLKCP", "SYNCS", "UCGABAR", "LDG", "STG", "LDS", "STS", "LDGSTS"]


def demangle(names):
    out = subprocess.run(["c++filt"], input="\n".join(names), capture_output=True, text=True).stdout.split("\n")
    return dict(zip(names, out))


def short(sig):
    s = sig.split("(")[0].replace("fdtd_b200::", "").replace("unsigned char", "u8").replace("unsigned short", "u16")
    return re.sub(r"\((int|bool)\)", "", s)


def main():
    want = sys.argv[1:]
    log = open(f"{ROOT}/fdtd_b200/csrc/build.log").read()
    res = {}
    for b in re.split(r"ptxas info\s+: Compiling entry function '", log)[1:]:
        name = b.split("'")[0]
        regs = re.search(r"Used (\d+) registers", b)
        sp = re.search(r"(\d+) bytes stack frame, (\d+) bytes spill stores, (\d+) bytes spill loads", b)
        sm = re.search(r"(\d+) bytes smem", b)
        res[name] = (regs.group(1) if regs else "?", sp.groups() if sp else ("?", "?", "?"), sm.group(1) if sm else "0")
    sass = subprocess.run(["cuobjdump", "-sass", f"{ROOT}/fdtd_b200/libfdtd_b200.so"], capture_output=True, text=True).stdout
    funcs = {}
    cur = None
    for line in sass.split("\n"):
        m = re.match(r"\s*Function : (\S+)", line)
        if m:
            cur = m.group(1); funcs[cur] = {"n": 0, **{k: 0 for k in MN}}; continue
        m = re.match(r"\s*/\*[0-9a-f]{4}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
        if cur and m:
            op = m.group(1).split(".")[0]
            if op.startswith("UCGABAR"): op = "UCGABAR"   # UCGABAR_ARV / UCGABAR_WAIT: barrier.cluster.arrive / wait
            funcs[cur]["n"] += 1
            if op in funcs[cur]: funcs[cur][op] += 1
    dm = demangle(list(funcs))
    rows = sorted((short(dm[f]), f) for f in funcs)
    print("| kernel | regs | stack | spill st/ld | static smem | SASS instr | " + " | ".join(MN) + " |")
    print("|---|" + "---|" * (5 + len(MN)))
    tot = {k: 0 for k in MN}
    for name, f in rows:
        for k in MN: tot[k] += funcs[f][k]
        if want and not any(w in name for w in want): continue
        r = res.get(f, ("?", ("?", "?", "?"), "0"))
        print(f"| `{name}` | {r[0]} | {r[1][0]} | {r[1][1]}/{r[1][2]} | {r[2]} | {funcs[f]['n']} | " + " | ".join(str(funcs[f][k]) for k in MN) + " |")
    print(f"\nWhole library: {len(funcs)} kernels, " + ", ".join(f"{tot[k]} × {k}" for k in ("UTMALDG", "UBLKCP", "SYNCS", "UCGABAR")) + ".")
    spills = sorted(short(dm[f]) for f in funcs if f in res and res[f][1][1] not in ("0", "?"))
    print(f"Kernels with spills ({len(spills)}): " + ("; ".join(f"`{s}`" for s in spills) if spills else "none") + ".")


if __name__ == "__main__":
    main()
