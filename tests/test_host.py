"""CPU tests of the host-side mirror of the reference interface (fdtd_b200/solver.py, driver.py, configs.py)."""
import ctypes as C
import math

import numpy as np
import pytest


def test_lattice_size_is_ceil_max_1(b200):
    # fdtd/mod.rs:25-29
    cfg = b200.FdtdSolverConfig(b200.Resolution((0.01, 0.01, 0.01), 0.001), b200.PhysicalConstants.REDUCED, (1.0, 1.0, 0.0))
    assert cfg.lattice_size() == (100, 100, 1) and cfg.num_cells() == 10000          # the app's own 100x100x1 lattice
    cfg = b200.FdtdSolverConfig(b200.Resolution((0.3, 0.3, 0.3), 0.1), b200.PhysicalConstants.SI, (1.0, 0.9, 0.29))
    assert cfg.lattice_size() == (4, 3, 1)
    cfg = b200.FdtdSolverConfig.from_lattice((512, 512, 512), b200.Resolution((1, 1, 1), 0.25), b200.PhysicalConstants.REDUCED)
    assert cfg.lattice_size() == (512, 512, 512)


def test_physical_constants_and_resolution_estimates(b200):
    si = b200.PhysicalConstants.SI
    assert abs(si.speed_of_light() - 299792458.0) < 1.0                                # material.rs:43-55
    assert b200.PhysicalConstants.REDUCED.speed_of_light() == 1.0
    r = b200.Resolution.estimate_from_min_wavelength(b200.PhysicalConstants.REDUCED, 9.0 * math.sqrt(3.0))
    assert abs(r.spatial[0] - 1.0) < 1e-15 and abs(r.temporal - 1.0 / math.sqrt(3.0)) < 1e-15   # Courant limit, mod.rs:40-45


def test_source_functions_match_oracle(b200, oracle):
    g = b200.GaussianPulse(12.5, 2.5)
    cw = b200.ContinousWave(0.3, 0.05)
    for t in np.linspace(0.0, 40.0, 97):
        assert g.evaluate(float(t)) == oracle.gaussian_pulse(float(t), 12.5, 2.5)       # source.rs:39-56
        assert cw.evaluate(float(t)) == oracle.continous_wave(float(t), 0.3, 0.05)      # source.rs:59-76
    v = g.with_amplitudes((0, 0, 50.0), (1.0, 0, 0)).evaluate(12.5)
    assert v.j == (0.0, 0.0, 50.0) and v.m == (1.0, 0.0, 0.0)


def test_normalize_point_bounds(b200):
    n = b200.normalize_point_bounds
    assert n(None, (4, 5, 6)) == ((0, 0, 0), (4, 5, 6))
    assert n(((1, 2, 3), (3, 4, 5)), (9, 9, 9)) == ((1, 2, 3), (3, 4, 5))
    assert n(("incl", (1, 2, 3), (2, 3, 4)), (9, 9, 9)) == ((1, 2, 3), (3, 4, 5))        # util.rs:151-152
    assert n(((3, 3, 3), (1, 5, 1)), (9, 9, 9)) == ((3, 3, 3), (3, 5, 3))               # util.rs:157-160


def test_sources_table_accumulates_time_like_the_reference(b200):
    from fdtd_b200 import driver

    src = b200.ContinousWave(0.0, 0.05).with_amplitudes((0, 0, 50.0), (0, 0, 0))
    s = driver.Sources([((1, 1, 1), src)])
    tab = s.table(0.0, 0.1, 30)
    t = 0.0
    for k in range(30):
        assert tab[k, 0, 2] == 50.0 * math.cos(math.tau * 0.05 * t)
        t += 0.1                                                                        # cpu/mod.rs:513
    assert tab.shape == (30, 1, 6) and not np.any(tab[:, :, 3:])


def test_stop_conditions(b200):
    from fdtd_b200 import driver

    class S:
        def __init__(self, tick, time):
            self._t, self._time = tick, time

        def tick(self):
            return self._t

        def time(self):
            return self._time

    assert not driver.StopCondition("never").reached(1e9, S(10**9, 1e9))
    assert driver.StopCondition("step_limit", 1000).reached(0, S(1000, 0)) and not driver.StopCondition("step_limit", 1000).reached(0, S(999, 0))
    assert driver.StopCondition("simulated_time_limit", 2.5).reached(0, S(0, 2.5))
    assert driver.StopCondition("realtime_limit", 3.0).reached(3.1, S(0, 0)) and not driver.StopCondition("realtime_limit", 3.0).reached(2.9, S(0, 0))


def test_scene_canonical_ids_match_library(b200, oracle):
    """configs.canonicalize (numpy) == fdtd_b200_canonicalize_material_index (C ABI) == oracle per-cell indexing."""
    from fdtd_b200 import _lib as L
    from fdtd_b200 import configs

    w = configs.config2_inclusions(48, 10)
    rows = w.palette_rows()
    ref_ids, ref_pal = oracle.index_materials(rows[w.ids])
    assert np.array_equal(ref_ids, w.ids) and np.array_equal(ref_pal, rows)             # already canonical
    # scramble the labelling and let the library canonicalise it
    perm = np.array([3, 0, 4, 1, 2])
    lab = perm[w.ids].astype(np.uint32)
    pal_in = np.zeros((5, 4)); pal_in[perm] = rows
    ids = np.zeros_like(lab)
    pal_out = (L.Material * 8)()
    n = C.c_uint32()
    rc = b200.lib().fdtd_b200_canonicalize_material_index(lab.ctypes.data_as(C.POINTER(C.c_uint32)), lab.size, np.ascontiguousarray(pal_in).ctypes.data_as(C.POINTER(L.Material)), 5,
                                                          ids.ctypes.data_as(C.POINTER(C.c_uint32)), pal_out, 8, C.byref(n))
    assert rc == 0 and n.value == 5 and np.array_equal(ids, w.ids)


def test_workloads_are_well_formed(b200):
    from fdtd_b200 import configs

    w1 = configs.config1_vacuum64()
    assert w1.lattice == (64, 64, 64) and w1.steps == 1000 and w1.source_points() == [(32, 32, 32)]
    w3 = configs.config3_heterogeneous(64, 10)
    assert w3.ids[0] == 0 and w3.ids.max() < len(w3.palette) and len(w3.palette) <= 1 + 16 + 4
    # the absorbing layer is impedance matched: sigma_m / mu = sigma / eps
    for m in w3.palette:
        if m.relative_permittivity == 1.0 and m.relative_permeability == 1.0 and m.eletrical_conductivity > 0 and m.magnetic_conductivity > 0:
            assert m.magnetic_conductivity == m.eletrical_conductivity
    a = configs.heterogeneous_slab_labels((32, 32, 48), 0, 48, n_balls=8)
    b = np.concatenate([configs.heterogeneous_slab_labels((32, 32, 48), 0, 20, n_balls=8), configs.heterogeneous_slab_labels((32, 32, 48), 20, 48, n_balls=8)])
    assert np.array_equal(a, b)                                                          # slabs tile the lattice


def test_per_cell_domain_description_is_indexed_by_first_occurrence(b200):
    class Checker(b200.DomainDescription):
        def material(self, p):
            return b200.Material(relative_permittivity=1.0 + ((p[0] + p[1] + p[2]) % 3))

    cfg = b200.FdtdSolverConfig.from_lattice((4, 3, 2), b200.Resolution((1, 1, 1), 0.25), b200.PhysicalConstants.REDUCED)
    inst = b200.FdtdB200Backend().create_instance(cfg, Checker())
    assert [m.relative_permittivity for m in inst.palette] == [1.0, 2.0, 3.0]
    exp = [(x + y + z) % 3 for z in range(2) for y in range(3) for x in range(4)]
    assert inst.ids.tolist() == exp
