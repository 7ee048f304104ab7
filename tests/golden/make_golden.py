"""Generates tests/golden/*.npz from the f64 oracle (oracle/fdtd_ref.hpp, T=double ≙ the reference CPU backend).

The reference ships no FDTD golden vectors and cannot be built here (Rust), so these
known-answer tests pin the ORACLE against regressions; they are small enough to be checked by hand against
cem-solver/src/fdtd/cpu/mod.rs:416-515.  Run:  python tests/golden/make_golden.py
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import oracle  # noqa: E402

OUT = os.path.dirname(os.path.abspath(__file__))
VAC = [[1.0, 0.0, 1.0, 0.0]]


def kat1_impulse():
    """8x8x8 vacuum, single j_z impulse at (4,4,4) during step 0, 4 steps: full E and H."""
    s = oracle.RefSolver((8, 8, 8), (1, 1, 1), 0.25)
    s.set_materials(VAC)
    jm = np.zeros((4, 1, 6))
    jm[0, 0, 2] = 1.0
    s.run(4, [(4, 4, 4)], jm)
    np.savez_compressed(os.path.join(OUT, "kat1_impulse.npz"), E=s.field(oracle.E), H=s.field(oracle.H))


def kat2_degenerate():
    """1x1xN, Nx1x1 and the app's own NxNx1 shape: derivatives along unit axes vanish (Neumann rule)."""
    out = {}
    for name, size, src in (("z", (1, 1, 16), (0, 0, 8)), ("x", (16, 1, 1), (8, 0, 0)), ("xy", (12, 12, 1), (6, 6, 0))):
        s = oracle.RefSolver(size, (1, 1, 1), 0.25)
        s.set_materials(VAC)
        jm = np.zeros((6, 1, 6))
        jm[:3, 0, :] = [1.0, -2.0, 3.0, 0.5, 0.25, -0.125]
        s.run(6, [src], jm)
        out[f"E_{name}"] = s.field(oracle.E)
        out[f"H_{name}"] = s.field(oracle.H)
    np.savez_compressed(os.path.join(OUT, "kat2_degenerate.npz"), **out)


def kat3_lossy_decay():
    """Uniform sigma=10, dt=0.001, REDUCED: a uniform E (zero curl) decays by exactly c_a per step."""
    s = oracle.RefSolver((4, 4, 4), (0.01, 0.01, 0.01), 0.001)
    s.set_materials([[1.0, 0.0, 1.0, 10.0]])
    e0 = np.ones((4, 4, 4, 3))
    s.write_field(oracle.E, e0)
    s.run(5)
    np.savez_compressed(os.path.join(OUT, "kat3_lossy_decay.npz"), E=s.field(oracle.E), H=s.field(oracle.H),
                        c=oracle.update_coefficients(0.001, 1.0, 1.0, [1.0, 0.0, 1.0, 10.0]))


def kat4_last_wins():
    """Two set_forcing calls on one cell in one step: the second wins (cpu/mod.rs:399-408)."""
    s = oracle.RefSolver((6, 6, 6), (1, 1, 1), 0.25)
    s.set_materials(VAC)
    s.begin_update()
    s.set_forcing((3, 3, 3), (1, 1, 1), (0, 0, 0))
    s.set_forcing((2, 3, 3), (0, 5, 0), (0, 0, 1))
    s.set_forcing((3, 3, 3), (0, 0, 7), (2, 0, 0))
    s.finish()
    np.savez_compressed(os.path.join(OUT, "kat4_last_wins.npz"), E=s.field(oracle.E), H=s.field(oracle.H))


def kat5_material_index():
    """First-occurrence ids over a 3-material checker."""
    mats = np.array([[1, 0, 2.5, 0], [1, 0, 1, 0], [2, 0.5, 1, 0.0]], dtype=np.float64)
    lab = np.array([(x + 2 * y + z) % 3 for z in range(3) for y in range(4) for x in range(5)])
    lab[0] = 1  # the first cell's material becomes id 0
    ids, pal = oracle.index_materials(mats[lab])
    np.savez_compressed(os.path.join(OUT, "kat5_material_index.npz"), labels=lab, materials=mats, ids=ids, palette=pal)


def kat6_projection():
    """Projection of a known field, identity transform: CPU colour path, fresh and stale buffer."""
    s = oracle.RefSolver((8, 6, 2), (1, 1, 1), 0.25)
    s.set_materials(VAC)
    jm = np.zeros((3, 1, 6))
    jm[:, 0, 2] = [0.5, 1.0, -0.5]
    s.run(3, [(4, 3, 0)], jm)
    eye = np.eye(4, dtype=np.float32)
    cmap = np.array([[0, 0, 4, 0.5], [0, 0, -4, 0.5], [1, 0, 0, 0.2], [0, 0, 0, 1]], dtype=np.float32)
    out = {}
    for mode in (0, 1, 2):
        out[f"fresh_{mode}"] = s.project(oracle.E, eye, cmap, mode, 8, 6, stale=False)
        out[f"stale_{mode}"] = s.project(oracle.E, eye, cmap, mode, 8, 6, stale=True)
    np.savez_compressed(os.path.join(OUT, "kat6_projection.npz"), cmap=cmap, **out)


if __name__ == "__main__":
    oracle.build()
    for fn in (kat1_impulse, kat2_degenerate, kat3_lossy_decay, kat4_last_wins, kat5_material_index, kat6_projection):
        fn()
        print("wrote", fn.__name__)
