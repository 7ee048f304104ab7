"""Fused step kernel against the number of distinct materials (coefficient table in shared memory or read through L1):
    python tools/many_materials.py [nx,ny,nz] [steps]
For each table size K the 512^3 lattice is filled (a) with ids drawn per cell and (b) with one id per z-plane, and the
kernel is timed with the table staged in shared memory up to 64 rows (default), 256 and 1024 (FDTD_B200_TUNE field 6)."""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402

import fdtd_b200 as b200  # noqa: E402
from fdtd_b200 import configs  # noqa: E402


def main():
    lattice = tuple(int(v) for v in (sys.argv[1] if len(sys.argv) > 1 else "512,512,512").split(","))
    steps = int(sys.argv[2]) if len(sys.argv) > 2 else 30
    nx, ny, nz = lattice
    rng = np.random.default_rng(7)
    for K in (21, 65, 200, 256, 300, 1000, 2000):
        pal = [b200.Material.VACUUM] + [b200.Material(relative_permittivity=1.0 + 0.01 * k, eletrical_conductivity=0.001 * (k % 7)) for k in range(1, K)]
        for layout in ("per-cell", "per-plane"):
            if layout == "per-cell":
                ids = rng.integers(0, K, nx * ny * nz, dtype=np.uint32)
            else:
                ids = np.repeat((np.arange(nz, dtype=np.uint32) * 7) % K, nx * ny)
            ids[:K] = np.arange(K, dtype=np.uint32)      # every material occurs, in palette order
            src = b200.GaussianPulse(12.5, 2.5).with_amplitudes((0.0, 0.0, 50.0), (0.0, 0.0, 0.0))
            w = configs.Workload("many", lattice, b200.Resolution((1.0, 1.0, 1.0), 0.25), b200.PhysicalConstants.REDUCED, pal, ids,
                                 [((nx // 2, ny // 2, nz // 2), src)], [], 0)
            for limit in (0, 256, 1024):
                os.environ["FDTD_B200_TUNE"] = f"0,0,0,-1,0,{limit}"
                inst = b200.FdtdB200Backend(fused_step=True).create_instance(w.solver_config(), w.domain())
                st = inst.create_state()
                st.set_source_table(w.source_points(), w.source_table(steps + 8))
                st.step(5)
                st.timer_start()
                st.step(steps)
                ms = st.timer_stop()
                print(json.dumps({"materials": K, "layout": layout, "table_smem_limit": limit or 64, "fused": bool(st.uses_fused),
                                  "gcells": w.n_cells * steps / (ms * 1e-3) / 1e9}), flush=True)
                st.close()


if __name__ == "__main__":
    main()
