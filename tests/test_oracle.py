"""CPU tests of the oracle (the C++ restatement of the reference): the reference's own two tests, the golden
known-answer vectors, analytic checks and an independent numpy formulation."""
import os

import numpy as np
import pytest

from helpers import rel_l2
from np_slab import NumpySlab

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
VAC = [[1.0, 0.0, 1.0, 0.0]]


def gold(name):
    return np.load(os.path.join(GOLD, name))


# ---- the reference's own tests: cem-solver/src/fdtd/util.rs:217-257 ---------------------------------
EXPECTED_POINTS = [[1, 2, 3], [2, 2, 3], [1, 3, 3], [2, 3, 3], [1, 2, 4], [2, 2, 4], [1, 3, 4], [2, 3, 4]]


def test_it_iters_inclusive(oracle):
    # x0..=x1 → end bound included → half-open end = x1 + 1 (util.rs:151-152)
    pts = oracle.iter_points((1, 2, 3), (2 + 1, 3 + 1, 4 + 1))
    assert pts.tolist() == EXPECTED_POINTS


def test_it_iters_exclusive(oracle):
    pts = oracle.iter_points((1, 2, 3), (3, 4, 5))
    assert pts.tolist() == EXPECTED_POINTS


def test_iter_points_empty_and_clamped(oracle):
    assert oracle.iter_points((2, 2, 2), (2, 2, 2)).shape == (0, 3)
    # end < start is clamped to start per axis (util.rs:157-160)
    assert oracle.iter_points((2, 2, 2), (1, 1, 1)).shape == (0, 3)


def test_strider_roundtrip(oracle):
    size = (5, 3, 4)
    for idx in (0, 1, 4, 5, 14, 15, 59):
        x, y, z = oracle.strider_point(size, idx)
        assert x + 5 * (y + 3 * z) == idx  # strider.rs:47-53


# ---- coefficients: fdtd/util.rs:99-128 ---------------------------------------------------------------
def test_update_coefficients_formula(oracle):
    dt, eps0, mu0 = 0.25, 2.0, 3.0
    mu_r, sm, eps_r, s = 1.5, 0.3, 2.5, 0.7
    c = oracle.update_coefficients(dt, eps0, mu0, [mu_r, sm, eps_r, s])
    xe = 0.5 * s * dt / (eps_r * eps0)
    xm = 0.5 * sm * dt / (mu_r * mu0)
    exp = [(1 - xe) / (1 + xe), dt / (eps_r * eps0 * (1 + xe)), (1 - xm) / (1 + xm), dt / (mu_r * mu0 * (1 + xm))]
    assert c.tolist() == exp


def test_vacuum_coefficients(oracle):
    assert oracle.update_coefficients(0.25, 1.0, 1.0, VAC[0]).tolist() == [1.0, 0.25, 1.0, 0.25]


# ---- golden known-answer vectors ---------------------------------------------------------------------
def test_kat1_impulse(oracle):
    g = gold("kat1_impulse.npz")
    s = oracle.RefSolver((8, 8, 8), (1, 1, 1), 0.25)
    s.set_materials(VAC)
    jm = np.zeros((4, 1, 6))
    jm[0, 0, 2] = 1.0
    s.run(4, [(4, 4, 4)], jm)
    assert np.array_equal(s.field(oracle.E), g["E"]) and np.array_equal(s.field(oracle.H), g["H"])
    # hand check of step 1: E_z(4,4,4) = c_a*0 + c_b*(0 - 1) = -0.25 and nothing else moved
    s1 = oracle.RefSolver((8, 8, 8), (1, 1, 1), 0.25)
    s1.set_materials(VAC)
    s1.run(1, [(4, 4, 4)], jm[:1])
    e = s1.field(oracle.E)
    assert e[4, 4, 4, 2] == -0.25 and np.count_nonzero(e) == 1 and np.count_nonzero(s1.field(oracle.H)) == 0
    # step 2: H picks up backward differences of E_z: Hx(4,4,4) = -d_b*(dEz/dy) = -0.25*(-0.25) and Hx(4,5,4) opposite
    s1.run(1)
    h = s1.field(oracle.H)
    assert h[4, 4, 4, 0] == 0.0625 and h[4, 5, 4, 0] == -0.0625
    assert h[4, 4, 4, 1] == -0.0625 and h[4, 4, 5, 1] == 0.0625


def test_kat2_degenerate(oracle):
    g = gold("kat2_degenerate.npz")
    for name, size, src in (("z", (1, 1, 16), (0, 0, 8)), ("x", (16, 1, 1), (8, 0, 0)), ("xy", (12, 12, 1), (6, 6, 0))):
        s = oracle.RefSolver(size, (1, 1, 1), 0.25)
        s.set_materials(VAC)
        jm = np.zeros((6, 1, 6))
        jm[:3, 0, :] = [1.0, -2.0, 3.0, 0.5, 0.25, -0.125]
        s.run(6, [src], jm)
        assert np.array_equal(s.field(oracle.E), g[f"E_{name}"])
        assert np.array_equal(s.field(oracle.H), g[f"H_{name}"])
    # in a 1x1xN lattice only d/dz survives: E_z and H_z never couple to anything
    ez, hz = g["E_z"][..., 2], g["H_z"][..., 2]
    assert ez[8, 0, 0] == -0.25 * 3.0 * 3 and np.count_nonzero(ez) == 1       # E_z' = E_z - c_b*j_z, three forced steps
    assert hz[8, 0, 0] == -0.25 * -0.125 * 3 and np.count_nonzero(hz) == 1    # H_z' = H_z - d_b*m_z


def test_kat3_lossy_decay(oracle):
    g = gold("kat3_lossy_decay.npz")
    s = oracle.RefSolver((4, 4, 4), (0.01, 0.01, 0.01), 0.001)
    s.set_materials([[1.0, 0.0, 1.0, 10.0]])
    s.write_field(oracle.E, np.ones((4, 4, 4, 3)))
    s.run(5)
    assert np.array_equal(s.field(oracle.E), g["E"])
    c_a = g["c"][0]
    assert abs(c_a - (1 - 0.005) / (1 + 0.005)) < 1e-16
    # uniform field has zero curl: |E| decays by c_a per step, H stays zero
    assert np.allclose(s.field(oracle.E), c_a ** 5, rtol=1e-14)
    assert np.count_nonzero(s.field(oracle.H)) == 0


def test_kat4_last_wins(oracle):
    g = gold("kat4_last_wins.npz")
    s = oracle.RefSolver((6, 6, 6), (1, 1, 1), 0.25)
    s.set_materials(VAC)
    s.begin_update()
    s.set_forcing((3, 3, 3), (1, 1, 1), (0, 0, 0))
    s.set_forcing((2, 3, 3), (0, 5, 0), (0, 0, 1))
    s.set_forcing((3, 3, 3), (0, 0, 7), (2, 0, 0))
    s.finish()
    e, h = s.field(oracle.E), s.field(oracle.H)
    assert np.array_equal(e, g["E"]) and np.array_equal(h, g["H"])
    assert h[3, 3, 3, 0] == -0.5 and h[3, 3, 2, 2] == -0.25            # m of the LAST registration
    assert e[3, 3, 3, 0] == 0.0                                        # j=(1,1,1) was overwritten
    with pytest.raises(IndexError):                                     # reference panics (cpu/mod.rs:392-396)
        s.begin_update()
        s.set_forcing((6, 0, 0), (0, 0, 0), (0, 0, 0))


def test_kat5_material_index(oracle):
    g = gold("kat5_material_index.npz")
    ids, pal = oracle.index_materials(g["materials"][g["labels"]])
    assert np.array_equal(ids, g["ids"]) and np.array_equal(pal, g["palette"])
    assert ids[0] == 0 and np.array_equal(pal[0], g["materials"][1])    # first cell's material is id 0
    # ids are handed out in first-occurrence order
    first = [int(np.argmax(ids == k)) for k in range(len(pal))]
    assert first == sorted(first)
    assert np.array_equal(pal[ids], g["materials"][g["labels"]])


def test_kat6_projection(oracle):
    g = gold("kat6_projection.npz")
    s = oracle.RefSolver((8, 6, 2), (1, 1, 1), 0.25)
    s.set_materials(VAC)
    jm = np.zeros((3, 1, 6))
    jm[:, 0, 2] = [0.5, 1.0, -0.5]
    s.run(3, [(4, 3, 0)], jm)
    eye = np.eye(4, dtype=np.float32)
    for mode in (0, 1, 2):
        assert np.array_equal(s.project(oracle.E, eye, g["cmap"], mode, 8, 6, stale=False), g[f"fresh_{mode}"])
        assert np.array_equal(s.project(oracle.E, eye, g["cmap"], mode, 8, 6, stale=True), g[f"stale_{mode}"])
    # the reference's projection shows the PREVIOUS step: stale(tick 3) == fresh(tick 2)
    s2 = oracle.RefSolver((8, 6, 2), (1, 1, 1), 0.25)
    s2.set_materials(VAC)
    s2.run(2, [(4, 3, 0)], jm[:2])
    assert np.array_equal(s2.project(oracle.E, eye, g["cmap"], 0, 8, 6, stale=False), g["stale_0"])


# ---- independent numpy formulation agrees with the C++ oracle ----------------------------------------
@pytest.mark.parametrize("size", [(7, 5, 6), (1, 1, 9), (10, 10, 1), (4, 1, 3)])
def test_numpy_formulation_matches_oracle(oracle, size):
    rng = np.random.default_rng(3)
    nx, ny, nz = size
    pal = np.array([[1, 0, 1, 0], [1.0, 0.0, 3.9, 0.0], [2.0, 0.5, 1.0, 0.0], [1.0, 0.0, 1.0, 10.0]])
    ids = rng.integers(0, 4, nx * ny * nz).astype(np.uint32)
    dx, dt = (1.0, 0.5, 2.0), 0.2
    s = oracle.RefSolver(size, dx, dt)
    s.set_materials(pal, ids, 1.0, 1.0)
    e0 = rng.standard_normal((nz, ny, nx, 3))
    h0 = rng.standard_normal((nz, ny, nx, 3))
    s.write_field(oracle.E, e0)
    s.write_field(oracle.H, h0)
    coef = np.array([oracle.update_coefficients(dt, 1.0, 1.0, pal[k]) for k in ids]).reshape(nz, ny, nx, 4)
    ns = NumpySlab(size, dx, dt, coef)
    for k, c in enumerate(("ex", "ey", "ez")):
        ns.f[c][1:-1] = e0[..., k]
    for k, c in enumerate(("hx", "hy", "hz")):
        ns.f[c][1:-1] = h0[..., k]
    src = (nx // 2, ny // 2, nz // 2)
    for step in range(5):
        j, m = (0.1 * step, -1.0, 2.0), (0.5, 0.0, -0.25 * step)
        s.begin_update(); s.set_forcing(src, j, m); s.finish()
        ns.set_forcing(src, j, m); ns.step()
    assert rel_l2(ns.field("E"), s.field(oracle.E)) < 1e-14
    assert rel_l2(ns.field("H"), s.field(oracle.H)) < 1e-14


def test_f32_oracle_tracks_f64(oracle):
    from fdtd_b200 import configs

    w = configs.config1_vacuum64(24, 60)
    from helpers import oracle_run

    d, _ = oracle_run(oracle, w, 60, oracle.F64)
    f, _ = oracle_run(oracle, w, 60, oracle.F32)
    assert rel_l2(f.field(oracle.E), d.field(oracle.E)) < 1e-5
    assert rel_l2(f.field(oracle.H), d.field(oracle.H)) < 1e-5
    assert d.tick == 60 and d.time == 15.0


def test_time_is_accumulated_not_multiplied(oracle):
    s = oracle.RefSolver((2, 2, 2), (1, 1, 1), 0.1)
    s.set_materials(VAC)
    s.run(7)
    t = 0.0
    for _ in range(7):
        t += 0.1
    assert s.time == t and s.time != 7 * 0.1  # cpu/mod.rs:513


def test_config2_full_fixture_is_what_the_oracle_produces(oracle):
    """tests/golden/config2_full.npz (tools/make_config2_golden.py, 30 min of host time) against a fresh oracle run of the
    first steps of the same 256³ workload: the fixture the GPU parity test trusts is the oracle's output."""
    import os

    from fdtd_b200 import configs

    g = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "config2_full.npz"))
    n_steps = 40     # the numerical light cone (one cell per step) reaches the probes 32 cells from the source
    w = configs.config2_inclusions(256, int(g["steps"]))
    s = oracle.RefSolver(w.lattice, w.resolution.spatial, w.resolution.temporal, precision=oracle.F32)
    s.set_materials(w.palette_rows(), w.ids, w.constants.vacuum_permittivity, w.constants.vacuum_permeability)
    series, _ = s.run(n_steps, w.source_points(), w.source_table(n_steps), [w.cell(p) for p in w.probes])
    assert np.array_equal(series.astype(np.float32), g["series_f32"][:n_steps])
    assert g["series_f32"].shape == (5000, 8, 6) and g["series_f64"].shape == (5000, 8, 6)
    assert np.any(g["series_f32"][:n_steps] != 0)          


def test_config3_full_fixture_is_what_the_oracle_produces(oracle):
    """tests/golden/config3_full.npz (tools/make_config3_golden.py: the 512³ bench workload, 200 steps) against a fresh oracle
    run of the same sources on the slab of planes around the source plane, for as many steps as the numerical light cone
    (one cell per step) needs to reach the slab's faces: inside it the probe series of the slab equal those of the lattice."""
    import os
    import sys

    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    sys.path.insert(0, root)
    from bench import build_workload
    from fdtd_b200 import configs

    g = np.load(os.path.join(root, "tests", "golden", "config3_full.npz"))
    assert g["series_f32"].shape == (200, 8, 6) and g["series_f64"].shape == (200, 8, 6) and int(g["steps"]) == 200
    lattice = tuple(int(v) for v in g["lattice"])
    half, n_steps = 20, 36     # a disturbance from the slab's faces needs 2 x 20 steps to reach the probe plane
    z0, z1 = lattice[2] // 2 - half, lattice[2] // 2 + half
    w = build_workload(lattice, z0, z1)
    sub = (lattice[0], lattice[1], z1 - z0)
    s = oracle.RefSolver(sub, w.resolution.spatial, w.resolution.temporal, precision=oracle.F32)
    s.set_materials(w.palette_rows(), w.ids, w.constants.vacuum_permittivity, w.constants.vacuum_permeability)
    pts = [(x, y, z - z0) for (x, y, z) in (p for p, _ in w.sources)]
    cells = [(z - z0) * sub[0] * sub[1] + y * sub[0] + x for (x, y, z) in w.probes]
    series, _ = s.run(n_steps, pts, w.source_table(n_steps), cells)
    assert np.array_equal(series.astype(np.float32), g["series_f32"][:n_steps])
    assert np.any(series[:, 0] != 0) and np.any(series[:, 1] != 0) and np.any(series[:, 4] != 0)   # probes 16 and 32 cells away
    assert np.array_equal(g["probes"], np.array(w.probes))
