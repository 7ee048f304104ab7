"""Generates tests/golden/config2_full.npz: BASELINE configs[1] at FULL size (256^3 lattice with dielectric / lossy /
magnetic inclusions and a conductor slab, CW source, 8 probes, 5000 steps) run through the CPU oracle (oracle/, the
restatement of cem-solver's CPU backend; the reference itself cannot be built in this environment).

Stored: the probe series of all 5000 steps in fp32-oracle arithmetic (≙ update.wgsl; the CUDA exact mode must reproduce
it bit for bit) and in f64 (≙ FdtdCpuBackend; north-star tolerance 1e-5), plus SHA-256 digests and norms of the full
E and H fields at steps 500 and 5000 (the fields themselves are 400 MB per snapshot).

Takes ≈ 30 min on 8 cores.  Run:  python tools/make_config2_golden.py
"""
import hashlib
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    import oracle
    from fdtd_b200 import configs

    n, steps, mid = 256, 5000, 500
    w = configs.config2_inclusions(n, steps)
    jm = w.source_table(steps)
    cells = [w.cell(p) for p in w.probes]
    out = {"lattice": np.array(w.lattice), "steps": np.array(steps), "mid": np.array(mid), "probes": np.array(w.probes)}
    for name, prec in (("f32", oracle.F32), ("f64", oracle.F64)):
        t0 = time.time()
        s = oracle.RefSolver(w.lattice, w.resolution.spatial, w.resolution.temporal, precision=prec, fast=False)
        s.set_materials(w.palette_rows(), w.ids, w.constants.vacuum_permittivity, w.constants.vacuum_permeability)
        a, _ = s.run(mid, w.source_points(), jm[:mid], cells)
        snaps = {}
        for tag in ("mid", "end"):
            if tag == "end":
                b, _ = s.run(steps - mid, w.source_points(), jm[mid:], cells)
            for comp, cname in ((oracle.E, "e"), (oracle.H, "h")):
                f = s.field(comp)
                snaps[f"{cname}_{tag}_norm"] = np.array(np.linalg.norm(f.ravel()))
                if prec == oracle.F32:
                    snaps[f"{cname}_{tag}_sha256"] = np.array(hashlib.sha256(np.ascontiguousarray(f.astype(np.float32)).tobytes()).hexdigest())
                else:   # a sparse sample of the f64 field for a tolerance check of the full field
                    snaps[f"{cname}_{tag}_sample"] = f[::8, ::8, ::8].copy()
        series = np.concatenate([a, b], axis=0)
        out[f"series_{name}"] = series.astype(np.float32) if prec == oracle.F32 else series
        for k, v in snaps.items():
            out[f"{name}_{k}"] = v
        assert s.tick == steps
        out[f"{name}_time"] = np.array(s.time)
        print(name, "done in", round(time.time() - t0), "s", flush=True)
        s.close()
    path = os.path.join(ROOT, "tests", "golden", "config2_full.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
