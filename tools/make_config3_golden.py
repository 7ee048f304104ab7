"""Generates tests/golden/config3_full.npz: the bench workload (BASELINE configs[2]: 512^3 heterogeneous lattice inside the
16-cell graded lossy layer, Gaussian point source, 8 probes — bench.build_workload) for 200 steps through the CPU oracle.
Stored: probe series in fp32-oracle and f64 arithmetic, SHA-256 + norms of the full fp32 fields at step 200 and a stride-5
sample of the f64 fields inside the box the pulse has reached (200 steps of dt = 0.25 carry it 50 cells), offset so that it
does not sit on the source cell's symmetry planes, where H cancels and a relative error means nothing.  ≈ 10 min on 8 cores, 18 GB of host memory for the f64 run.  Run: python tools/make_config3_golden.py"""
import hashlib
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


C, R, S = 256, 60, 5
SAMPLE = (slice(C - R, C + R, S), slice(C - R + 1, C + R, S), slice(C - R + 2, C + R, S))   # (z, y, x)


def main():
    import oracle
    from bench import build_workload

    lattice, steps = (512, 512, 512), 200
    w = build_workload(lattice, 0, lattice[2])
    jm = w.source_table(steps)
    cells = [w.cell(p) for p in w.probes]
    out = {"lattice": np.array(lattice), "steps": np.array(steps), "probes": np.array(w.probes),
           "sample_slices": np.array([[sl.start, sl.stop, sl.step] for sl in SAMPLE])}
    for name, prec in (("f32", oracle.F32), ("f64", oracle.F64)):
        t0 = time.time()
        s = oracle.RefSolver(w.lattice, w.resolution.spatial, w.resolution.temporal, precision=prec, fast=False)
        s.set_materials(w.palette_rows(), w.ids, w.constants.vacuum_permittivity, w.constants.vacuum_permeability)
        series, _ = s.run(steps, w.source_points(), jm, cells)
        for comp, cname in ((oracle.E, "e"), (oracle.H, "h")):
            f = s.field(comp)
            out[f"{name}_{cname}_norm"] = np.array(np.linalg.norm(f.ravel()))
            if prec == oracle.F32:
                out[f"{name}_{cname}_sha256"] = np.array(hashlib.sha256(np.ascontiguousarray(f.astype(np.float32)).tobytes()).hexdigest())
            else:
                out[f"{name}_{cname}_sample"] = f[SAMPLE].copy()
            del f
        out[f"series_{name}"] = series.astype(np.float32) if prec == oracle.F32 else series
        out[f"{name}_time"] = np.array(s.time)
        print(name, "done in", round(time.time() - t0), "s", flush=True)
        s.close()
    path = os.path.join(ROOT, "tests", "golden", "config3_full.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
