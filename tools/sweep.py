"""Launch-shape sweep of the float4 z-streaming kernels on the bench lattice (FDTD_B200_TUNE=bx,by,zc)."""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402

import fdtd_b200  # noqa: E402
from bench import build_workload  # noqa: E402


def main():
    lattice = tuple(int(v) for v in (sys.argv[1] if len(sys.argv) > 1 else "512,512,512").split(","))
    steps = int(sys.argv[2]) if len(sys.argv) > 2 else 10
    w = build_workload(lattice, 0, lattice[2])
    tab = w.source_table(steps + 4)
    rows = []
    shapes = [(128, 1), (128, 4)]
    zcs = [16]
    variants = [("fast", False, False), ("exact", True, False)]
    for name, exact, generic in variants:
        for bx, by in shapes:
            if bx * 4 > lattice[0] and bx != 32:
                continue
            for zc in zcs:
                if name == "exact" and zc not in (16, 32):
                    continue
                os.environ["FDTD_B200_TUNE"] = f"{bx},{by},{zc},0,0"
                inst = fdtd_b200.FdtdB200Backend(exact_arithmetic=exact).create_instance(w.solver_config(), w.domain())
                st = inst.create_state()
                st.set_source_table(w.source_points(), tab)
                st.step(3)
                st.set_profiling(True)
                st.timer_start()
                st.step(steps)
                ms = st.timer_stop()
                ph, pe = st.profile(0), st.profile(1)
                g = w.n_cells * steps / (ms * 1e-3) / 1e9
                rows.append({"mode": name, "bx": bx, "by": by, "zc": zc, "gcells": g, "gbs": 74 * g, "h_ms": ph[0] / ph[1], "e_ms": pe[0] / pe[1]})
                print(json.dumps(rows[-1]), flush=True)
                st.close()
    # TMA bulk-copy staged kernels: (by, stages) x zc
    for by, stg in [(1, 4), (2, 3)]:
        for zc in (32, 64, 128):
            for name, exact in (("bulk", False), ("bulk_exact", True)):
                os.environ["FDTD_B200_TUNE"] = f"128,{by},{zc},1,{stg}"
                inst = fdtd_b200.FdtdB200Backend(exact_arithmetic=exact).create_instance(w.solver_config(), w.domain())
                st = inst.create_state()
                st.set_source_table(w.source_points(), tab)
                st.step(3)
                st.set_profiling(True)
                st.timer_start()
                st.step(steps)
                ms = st.timer_stop()
                ph, pe = st.profile(0), st.profile(1)
                g = w.n_cells * steps / (ms * 1e-3) / 1e9
                rows.append({"mode": name, "bx": 128, "by": by, "stages": stg, "zc": zc, "gcells": g, "gbs": 74 * g, "h_ms": ph[0] / ph[1], "e_ms": pe[0] / pe[1]})
                print(json.dumps(rows[-1]), flush=True)
                st.close()
    os.environ["FDTD_B200_TUNE"] = ""
    if len(sys.argv) > 3 and sys.argv[3] == "nogeneric":
        print("BEST", json.dumps(max(rows, key=lambda r: r["gcells"])))
        return
    inst = fdtd_b200.FdtdB200Backend(force_generic_kernels=True).create_instance(w.solver_config(), w.domain())
    st = inst.create_state()
    st.set_source_table(w.source_points(), tab)
    st.step(2)
    st.timer_start()
    st.step(steps)
    ms = st.timer_stop()
    g = w.n_cells * steps / (ms * 1e-3) / 1e9
    print(json.dumps({"mode": "generic", "gcells": g, "gbs": 74 * g}), flush=True)
    best = max(rows, key=lambda r: r["gcells"])
    print("BEST", json.dumps(best))


if __name__ == "__main__":
    main()
