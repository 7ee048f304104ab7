"""GPU parity tests: the CUDA path, called through the C ABI (ctypes), against the oracle on identical lattices,
sources and step counts.

Bars (BASELINE.json north_star):
  * FDTD_B200_EXACT_ARITHMETIC vs oracle<float> (the WGSL arithmetic): BIT-IDENTICAL fields and probe series;
  * default (reciprocal multiply + FMA contraction) vs oracle<float>: rel-L2 <= 5e-6;
  * either vs oracle<double> (the reference CPU backend): rel-L2 <= 1e-5;
  * tick, material ids, coefficient table: exact.
"""
import os

import numpy as np
import pytest

from helpers import b200_run, b200_state, oracle_run, random_fields, rel_l2

pytestmark = pytest.mark.gpu

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
TOL_FAST_VS_F32 = 5e-6
TOL_VS_F64 = 1e-5


def _cfg(b200, size, dx=(1, 1, 1), dt=0.25):
    from fdtd_b200 import configs

    return configs.Workload("t", tuple(size), b200.Resolution(tuple(dx), dt), b200.PhysicalConstants.REDUCED,
                            [b200.Material.VACUUM], None, [], [], 0)


def _fields(st, b200):
    return st.read_field(b200.FieldComponent.E), st.read_field(b200.FieldComponent.H)


def _random_workload(b200, size, seed, dx=(1.0, 0.5, 2.0), dt=0.2, n_mat=4):
    from fdtd_b200 import configs

    rng = np.random.default_rng(seed)
    pal = [b200.Material.VACUUM, b200.Material(relative_permittivity=3.9), b200.Material(2.0, 0.5, 1.0, 0.0),
           b200.Material(eletrical_conductivity=10.0)][:n_mat]
    nx, ny, nz = size
    ids = rng.integers(0, len(pal), nx * ny * nz).astype(np.uint32)
    ids[0] = 0
    src_pt = (nx // 2, ny // 2, nz // 2)
    src = b200.ContinousWave(0.3, 0.07).with_amplitudes((1.0, -2.0, 3.0), (0.5, 0.25, -0.125))
    return configs.Workload("rnd", tuple(size), b200.Resolution(dx, dt), b200.PhysicalConstants.REDUCED, pal, ids,
                            [(src_pt, src)], [], 0)


# ---- bit-exact parity against the fp32 oracle ---------------------------------------------------------
@pytest.mark.parametrize("size,generic", [
    ((32, 16, 12), False), ((32, 16, 12), True), ((128, 8, 6), False), ((132, 5, 7), False), ((516, 3, 5), False),
    ((7, 5, 6), False), ((1, 1, 33), False), ((33, 1, 1), False), ((100, 100, 1), False), ((4, 4, 4), False),
    ((4, 1, 1), False), ((8, 1, 40), False),
])
def test_exact_mode_is_bit_identical_to_f32_oracle(b200, oracle, size, generic):
    w = _random_workload(b200, size, seed=sum(size))
    init = random_fields(w, seed=1)
    steps = 12
    ref, _ = oracle_run(oracle, w, steps, oracle.F32, init=init)
    _, st, _ = b200_run(b200, w, steps, exact=True, generic=generic, init=init)
    e, h = _fields(st, b200)
    assert st.uses_vec4 == (size[0] % 4 == 0 and not generic)
    assert st.tick() == steps == ref.tick and st.time() == ref.time
    assert np.array_equal(e, ref.field(oracle.E).astype(np.float32))
    assert np.array_equal(h, ref.field(oracle.H).astype(np.float32))


@pytest.mark.parametrize("size", [(32, 10, 9), (10, 9, 8)])
def test_exact_mode_with_true_division(b200, oracle, size):
    """dx not a power of two → the kernels divide (update.wgsl:159,171) instead of multiplying by 1/dx."""
    w = _random_workload(b200, size, seed=21, dx=(0.3, 0.7, 1.1), dt=0.11)
    init = random_fields(w, seed=6)
    ref, _ = oracle_run(oracle, w, 15, oracle.F32, init=init)
    _, st, _ = b200_run(b200, w, 15, exact=True, init=init)
    e, h = _fields(st, b200)
    assert np.array_equal(e, ref.field(oracle.E).astype(np.float32)) and np.array_equal(h, ref.field(oracle.H).astype(np.float32))
    _, stf, _ = b200_run(b200, w, 15, exact=False, init=init)
    ef, hf = _fields(stf, b200)
    assert rel_l2(ef, ref.field(oracle.E)) <= TOL_FAST_VS_F32 and rel_l2(hf, ref.field(oracle.H)) <= TOL_FAST_VS_F32


@pytest.mark.parametrize("size", [(64, 24, 20), (36, 9, 11), (9, 10, 11)])
def test_fast_mode_within_tolerance(b200, oracle, size):
    w = _random_workload(b200, size, seed=5)
    init = random_fields(w, seed=2)
    steps = 40
    ref32, _ = oracle_run(oracle, w, steps, oracle.F32, init=init)
    ref64, _ = oracle_run(oracle, w, steps, oracle.F64, init=init)
    _, st, _ = b200_run(b200, w, steps, exact=False, init=init)
    e, h = _fields(st, b200)
    assert rel_l2(e, ref32.field(oracle.E)) <= TOL_FAST_VS_F32 and rel_l2(h, ref32.field(oracle.H)) <= TOL_FAST_VS_F32
    assert rel_l2(e, ref64.field(oracle.E)) <= TOL_VS_F64 and rel_l2(h, ref64.field(oracle.H)) <= TOL_VS_F64


# ---- BASELINE config 1: 64^3 vacuum, Gaussian pulse, 1000 steps ---------------------------------------
def test_config1_vacuum64_1000_steps(b200, oracle):
    from fdtd_b200 import configs

    w = configs.config1_vacuum64()
    probes = [(32, 32, 32), (40, 32, 32), (32, 20, 50), (0, 0, 0), (63, 63, 63)]
    ref64, s64 = oracle_run(oracle, w, 1000, oracle.F64, probes=probes)
    ref32, s32 = oracle_run(oracle, w, 1000, oracle.F32, probes=probes)
    _, st_x, sx = b200_run(b200, w, 1000, exact=True, probes=probes)
    _, st_f, sf = b200_run(b200, w, 1000, exact=False, probes=probes)
    ex, hx = _fields(st_x, b200)
    ef, hf = _fields(st_f, b200)
    assert np.array_equal(ex, ref32.field(oracle.E).astype(np.float32)) and np.array_equal(hx, ref32.field(oracle.H).astype(np.float32))
    assert np.array_equal(sx, s32.astype(np.float32))                       # probe series bit-identical
    for got_e, got_h in ((ex, hx), (ef, hf)):
        assert rel_l2(got_e, ref64.field(oracle.E)) <= TOL_VS_F64
        assert rel_l2(got_h, ref64.field(oracle.H)) <= TOL_VS_F64
    assert rel_l2(sf, s64) <= TOL_VS_F64 and rel_l2(sx, s64) <= TOL_VS_F64
    assert st_x.tick() == 1000 and st_x.time() == ref64.time


def test_config1_intermediate_steps(b200, oracle):
    """Compare full E,H at steps {1, 10, 100} with one persistent handle."""
    from fdtd_b200 import configs

    w = configs.config1_vacuum64()
    ref = oracle.RefSolver(w.lattice, w.resolution.spatial, w.resolution.temporal, precision=oracle.F64)
    ref.set_materials(w.palette_rows())
    inst, st = b200_state(b200, w)
    tab = w.source_table(100)
    st.set_source_table(w.source_points(), tab)
    done = 0
    for upto in (1, 10, 100):
        ref.run(upto - done, w.source_points(), tab[done:upto])
        st.step(upto - done)
        done = upto
        assert rel_l2(st.read_field(b200.FieldComponent.E), ref.field(oracle.E)) <= TOL_VS_F64
        h = st.read_field(b200.FieldComponent.H)
        if np.any(ref.field(oracle.H)):
            assert rel_l2(h, ref.field(oracle.H)) <= TOL_VS_F64
        assert st.tick() == upto


# ---- BASELINE config 2 (scaled) : inclusions + probes -------------------------------------------------
def test_config2_inclusions_probes(b200, oracle):
    from fdtd_b200 import configs

    w = configs.config2_inclusions(96, 400)
    ref64, s64 = oracle_run(oracle, w, 400, oracle.F64, probes=w.probes)
    ref32, s32 = oracle_run(oracle, w, 400, oracle.F32, probes=w.probes)
    _, st, series = b200_run(b200, w, 400, exact=False, probes=w.probes)
    _, stx, series_x = b200_run(b200, w, 400, exact=True, probes=w.probes)
    assert np.array_equal(st.material_ids(), w.ids)                          # integer material indexing: bit-exact
    assert st.index_bytes == 1
    e, h = _fields(st, b200)
    assert rel_l2(e, ref64.field(oracle.E)) <= TOL_VS_F64 and rel_l2(h, ref64.field(oracle.H)) <= TOL_VS_F64
    assert rel_l2(series, s64) <= TOL_VS_F64
    for k in range(len(w.probes)):
        assert rel_l2(series[:, k], s64[:, k]) <= 5 * TOL_VS_F64            # every probe series individually
    assert np.array_equal(series_x, s32.astype(np.float32))
    assert np.array_equal(stx.read_field(b200.FieldComponent.E), ref32.field(oracle.E).astype(np.float32))


# ---- BASELINE config 3 (scaled): heterogeneous lattice + lossy absorbing layer ------------------------
def test_config3_heterogeneous_scaled(b200, oracle):
    from fdtd_b200 import configs

    w = configs.config3_heterogeneous(128, 100)
    ref64, _ = oracle_run(oracle, w, 100, oracle.F64)
    _, st, _ = b200_run(b200, w, 100)
    e, h = _fields(st, b200)
    assert rel_l2(e, ref64.field(oracle.E)) <= TOL_VS_F64 and rel_l2(h, ref64.field(oracle.H)) <= TOL_VS_F64
    tab = st.coefficient_table()
    exp = np.array([oracle.update_coefficients(0.25, 1.0, 1.0, m.as_row()) for m in w.palette]).astype(np.float32)
    assert np.array_equal(tab, exp)                                          # f64 derivation, one cast to f32


# ---- golden known-answer vectors through the C ABI -----------------------------------------------------
def test_golden_kat1_impulse(b200):
    from fdtd_b200 import configs

    g = np.load(os.path.join(GOLD, "kat1_impulse.npz"))
    w = _cfg(b200, (8, 8, 8))
    inst, st = b200_state(b200, w, exact=True)
    jm = np.zeros((4, 1, 6), dtype=np.float32)
    jm[0, 0, 2] = 1.0
    st.set_source_table([(4, 4, 4)], jm)
    st.step(4)
    # every value in this KAT is a short dyadic rational: fp32 reproduces the f64 golden vector exactly
    assert np.array_equal(st.read_field(b200.FieldComponent.E).astype(np.float64), g["E"])
    assert np.array_equal(st.read_field(b200.FieldComponent.H).astype(np.float64), g["H"])


def test_golden_kat2_degenerate(b200):
    g = np.load(os.path.join(GOLD, "kat2_degenerate.npz"))
    for name, size, src in (("z", (1, 1, 16), (0, 0, 8)), ("x", (16, 1, 1), (8, 0, 0)), ("xy", (12, 12, 1), (6, 6, 0))):
        w = _cfg(b200, size)
        inst, st = b200_state(b200, w, exact=True)
        jm = np.zeros((6, 1, 6), dtype=np.float32)
        jm[:3, 0, :] = [1.0, -2.0, 3.0, 0.5, 0.25, -0.125]
        st.set_source_table([src], jm)
        st.step(6)
        assert rel_l2(st.read_field(b200.FieldComponent.E), g[f"E_{name}"]) <= 1e-6
        assert rel_l2(st.read_field(b200.FieldComponent.H), g[f"H_{name}"]) <= 1e-6


def test_golden_kat3_lossy_decay(b200):
    from fdtd_b200 import configs

    g = np.load(os.path.join(GOLD, "kat3_lossy_decay.npz"))
    w = configs.Workload("lossy", (4, 4, 4), b200.Resolution((0.01, 0.01, 0.01), 0.001), b200.PhysicalConstants.REDUCED,
                         [b200.Material(eletrical_conductivity=10.0)], None, [], [], 0)
    inst, st = b200_state(b200, w)
    st.write_field(b200.FieldComponent.E, np.ones((4, 4, 4, 3), dtype=np.float32))
    st.step(5)
    assert rel_l2(st.read_field(b200.FieldComponent.E), g["E"]) <= 1e-6
    assert not np.any(st.read_field(b200.FieldComponent.H))


def test_golden_kat4_last_wins_through_update_pass(b200, oracle):
    g = np.load(os.path.join(GOLD, "kat4_last_wins.npz"))
    w = _cfg(b200, (6, 6, 6))
    inst, st = b200_state(b200, w, exact=True)
    up = inst.begin_update(st)
    up.set_forcing((3, 3, 3), b200.SourceValues((1, 1, 1), (0, 0, 0)))
    up.set_forcing((2, 3, 3), b200.SourceValues((0, 5, 0), (0, 0, 1)))
    up.set_forcing((3, 3, 3), b200.SourceValues((0, 0, 7), (2, 0, 0)))
    up.finish()
    assert np.array_equal(st.read_field(b200.FieldComponent.E).astype(np.float64), g["E"])
    assert np.array_equal(st.read_field(b200.FieldComponent.H).astype(np.float64), g["H"])
    with pytest.raises(IndexError):                      # the reference panics (cpu/mod.rs:392-396)
        inst.begin_update(st).set_forcing((6, 0, 0), b200.SourceValues())
    # forcing is per step: the next unforced step must not re-apply it (cpu/mod.rs:370-377)
    inst.begin_update(st).finish()
    ref = oracle.RefSolver((6, 6, 6), (1, 1, 1), 0.25, precision=oracle.F32)
    ref.set_materials([[1.0, 0.0, 1.0, 0.0]])
    ref.begin_update()
    ref.set_forcing((3, 3, 3), (1, 1, 1), (0, 0, 0))
    ref.set_forcing((2, 3, 3), (0, 5, 0), (0, 0, 1))
    ref.set_forcing((3, 3, 3), (0, 0, 7), (2, 0, 0))
    ref.finish()
    ref.begin_update()
    ref.finish()
    assert st.tick() == 2
    assert np.array_equal(st.read_field(b200.FieldComponent.E), ref.field(oracle.E).astype(np.float32))
    assert np.array_equal(st.read_field(b200.FieldComponent.H), ref.field(oracle.H).astype(np.float32))


def test_golden_kat6_projection(b200):
    g = np.load(os.path.join(GOLD, "kat6_projection.npz"))
    w = _cfg(b200, (8, 6, 2))
    inst, st = b200_state(b200, w, exact=True)
    jm = np.zeros((3, 1, 6), dtype=np.float32)
    jm[:, 0, 2] = [0.5, 1.0, -0.5]
    st.set_source_table([(4, 3, 0)], jm)
    eye = np.eye(4, dtype=np.float32)
    imgs = {}
    for tick in (2, 3):
        st.step(tick - st.tick())
        for mode in (0, 1, 2):
            target = b200.ImageTarget(8, 6)
            from fdtd_b200.solver import B200Projection

            proj = B200Projection(target, b200.ProjectionParameters(eye, b200.FieldComponent.E, g["cmap"]), mode=mode)
            p = inst.begin_projection_pass(st)
            p.add_projection(proj)
            p.finish()
            imgs[(tick, mode)] = target.image.copy()
    for mode in (0, 1, 2):
        # fresh field at tick 3; sRGB pow() may differ from the host libm by one LSB
        assert np.max(np.abs(imgs[(3, mode)].astype(int) - g[f"fresh_{mode}"].astype(int))) <= (1 if mode == 0 else 0)
        # the reference's picture at tick 3 is the previous step's field = our picture at tick 2
        assert np.max(np.abs(imgs[(2, mode)].astype(int) - g[f"stale_{mode}"].astype(int))) <= (1 if mode == 0 else 0)


# ---- reference-style driving (one update pass per step) == batched stepping ---------------------------
def test_reference_style_loop_equals_batched(b200):
    from fdtd_b200 import configs, driver

    w = configs.config1_vacuum64(32, 60)
    inst, st_a = b200_state(b200, w)
    srcs = driver.Sources(w.sources)
    n = driver.run_reference_style(inst, st_a, srcs, driver.StopCondition("step_limit", 60))
    assert n == 60 and st_a.tick() == 60
    _, st_b = b200_state(b200, w)
    driver.run_batched(st_b, srcs, 60, w.resolution.temporal, chunk=25)
    assert st_b.tick() == 60 and st_b.time() == st_a.time()
    assert np.array_equal(st_a.read_field(b200.FieldComponent.E), st_b.read_field(b200.FieldComponent.E))
    assert np.array_equal(st_a.read_field(b200.FieldComponent.H), st_b.read_field(b200.FieldComponent.H))


# ---- field views ---------------------------------------------------------------------------------------
def test_field_view_at_and_iter(b200, oracle):
    w = _random_workload(b200, (12, 7, 5), seed=9)
    init = random_fields(w, seed=3)
    inst, st = b200_state(b200, w)
    st.write_field(b200.FieldComponent.E, init[0])
    st.write_field(b200.FieldComponent.H, init[1])
    view = inst.field(st, ((2, 1, 1), (5, 4, 3)), b200.FieldComponent.H)
    pts = [p for p, _ in view.iter()]
    assert pts == [tuple(int(v) for v in p) for p in oracle.iter_points((2, 1, 1), (5, 4, 3))]  # x-fastest (util.rs:174-199)
    for p, v in view.iter():
        assert np.array_equal(v, init[1][p[2], p[1], p[0]].astype(np.float64))
    assert view.at((1, 1, 1)) is None and view.at((5, 1, 1)) is None                             # outside the range
    assert np.array_equal(view.at((4, 3, 2)), init[1][2, 3, 4].astype(np.float64))
    whole = inst.field(st, None, b200.FieldComponent.E).array()
    assert np.array_equal(whole, init[0].astype(np.float64))
    incl = inst.field(st, ("incl", (0, 0, 0), (1, 1, 1)), b200.FieldComponent.E).array()
    assert incl.shape == (2, 2, 2, 3)
    empty = inst.field(st, ((3, 3, 3), (3, 3, 3)), b200.FieldComponent.E).array()
    assert empty.size == 0


# ---- u16 material index ---------------------------------------------------------------------------------
def test_u16_material_index(b200, oracle):
    from fdtd_b200 import configs

    rng = np.random.default_rng(4)
    n_mat = 700
    pal = [b200.Material(1.0 + 0.001 * k, 0.0, 1.0 + 0.01 * k, 0.001 * (k % 7)) for k in range(n_mat)]
    size = (16, 12, 10)
    ids = rng.integers(0, n_mat, size[0] * size[1] * size[2]).astype(np.uint32)
    w = configs.Workload("u16", size, b200.Resolution((1, 1, 1), 0.25), b200.PhysicalConstants.REDUCED, pal, ids,
                         [((8, 6, 5), b200.GaussianPulse(2.0, 1.0).with_amplitudes((0, 0, 5.0), (0, 0, 0)))], [], 0)
    init = random_fields(w, seed=5)
    ref, _ = oracle_run(oracle, w, 10, oracle.F32, init=init)
    _, st, _ = b200_run(b200, w, 10, exact=True, init=init)
    assert st.index_bytes == 2 and np.array_equal(st.material_ids(), ids)
    assert np.array_equal(st.read_field(b200.FieldComponent.E), ref.field(oracle.E).astype(np.float32))
    assert np.array_equal(st.read_field(b200.FieldComponent.H), ref.field(oracle.H).astype(np.float32))


# ---- many sources / many probes (binary-search path of the fused lookup) -------------------------------
def test_many_sources_and_probes(b200, oracle):
    from fdtd_b200 import configs

    size = (24, 10, 9)
    rng = np.random.default_rng(6)
    cells = rng.choice(size[0] * size[1] * size[2], 150, replace=False)
    pts = [(int(c % 24), int((c // 24) % 10), int(c // 240)) for c in cells]
    srcs = [(p, b200.ContinousWave(0.1 * k, 0.03).with_amplitudes((1, 0.5 * k, -1), (0.2, 0, 0.1 * k))) for k, p in enumerate(pts)]
    w = configs.Workload("many", size, b200.Resolution((1, 1, 1), 0.25), b200.PhysicalConstants.REDUCED, [b200.Material.VACUUM], None, srcs, [], 0)
    probes = pts[:40] + pts[:3]                      # duplicates allowed
    ref, s32 = oracle_run(oracle, w, 15, oracle.F32, probes=probes)
    for generic in (False, True):
        _, st, series = b200_run(b200, w, 15, exact=True, generic=generic, probes=probes)
        assert np.array_equal(st.read_field(b200.FieldComponent.E), ref.field(oracle.E).astype(np.float32))
        assert np.array_equal(series, s32.astype(np.float32))


# ---- z-slab decomposition: two handles on one GPU == one handle (bit-identical) -------------------------
@pytest.mark.parametrize("size,cut", [((32, 12, 16), 8), ((32, 12, 16), 1), ((9, 7, 10), 6)])
def test_two_slabs_equal_one_domain(b200, size, cut):
    from fdtd_b200 import _lib as L

    w = _random_workload(b200, size, seed=11)
    init = random_fields(w, seed=4)
    steps = 9
    _, whole, _ = b200_run(b200, w, steps, init=init)
    nz = size[2]
    _, lo = b200_state(b200, w, z_slab=(0, cut))
    _, hi = b200_state(b200, w, z_slab=(cut, nz))
    lib = b200.lib()
    tab = w.source_table(steps)
    for st, (a, b) in ((lo, (0, cut)), (hi, (cut, nz))):
        st.write_field(b200.FieldComponent.E, init[0][a:b], (0, 0, a), (size[0], size[1], b))
        st.write_field(b200.FieldComponent.H, init[1][a:b], (0, 0, a), (size[0], size[1], b))
        st.set_source_table(w.source_points(), tab)

    def h_down():
        L.check(lib.fdtd_b200_halo_copy(lo._h, L.HALO_RECV_HX_HIGH, hi._h, L.HALO_SEND_HX_BOTTOM))
        L.check(lib.fdtd_b200_halo_copy(lo._h, L.HALO_RECV_HY_HIGH, hi._h, L.HALO_SEND_HY_BOTTOM))

    def e_up():
        L.check(lib.fdtd_b200_halo_copy(hi._h, L.HALO_RECV_EX_LOW, lo._h, L.HALO_SEND_EX_TOP))
        L.check(lib.fdtd_b200_halo_copy(hi._h, L.HALO_RECV_EY_LOW, lo._h, L.HALO_SEND_EY_TOP))

    h_down(); e_up()                                  # ghost planes of the initial condition
    for _ in range(steps):
        lo.update_h(0, cut); hi.update_h(0, nz - cut)
        h_down()
        lo.update_e(0, cut); hi.update_e(0, nz - cut)
        e_up()
        lo.end_step(); hi.end_step()
    for comp in (b200.FieldComponent.E, b200.FieldComponent.H):
        got = np.concatenate([lo.read_field(comp), hi.read_field(comp)], axis=0)
        assert np.array_equal(got, whole.read_field(comp))
    assert lo.tick() == steps == hi.tick()


# ---- size-independent properties at full benchmark size -------------------------------------------------
def test_full_size_linearity_and_zero_state(b200):
    """512^3 (the bench workload): (a) zero fields + no sources stay exactly zero; (b) the update is linear —
    doubling the source doubles every field value exactly (powers of two commute with fp32 rounding)."""
    from fdtd_b200 import configs

    w = configs.config3_heterogeneous(512, 8)
    inst, st = b200_state(b200, w)
    st.step(2)
    box = ((200, 200, 250), (328, 328, 262))
    assert not np.any(st.read_field(b200.FieldComponent.E, *box)) and not np.any(st.read_field(b200.FieldComponent.H, *box))
    assert np.array_equal(st.material_ids()[:: 4099], w.ids[:: 4099])
    st.reset()
    tab = w.source_table(8)
    st.set_source_table(w.source_points(), tab)
    st.step(8)
    e1 = st.read_field(b200.FieldComponent.E, *box)
    h1 = st.read_field(b200.FieldComponent.H, *box)
    assert np.any(e1) and np.any(h1) and np.all(np.isfinite(e1))
    st.reset()
    st.set_source_table(w.source_points(), 2.0 * tab)
    st.step(8)
    assert np.array_equal(st.read_field(b200.FieldComponent.E, *box), 2.0 * e1)
    assert np.array_equal(st.read_field(b200.FieldComponent.H, *box), 2.0 * h1)
    assert st.tick() == 8 and st.time() == 2.0


def test_full_size_matches_oracle_on_a_short_run(b200, oracle):
    """256^3 config-2 lattice, 30 steps against the f64 oracle (a few seconds of CPU)."""
    from fdtd_b200 import configs

    w = configs.config2_inclusions(256, 30)
    ref64, s64 = oracle_run(oracle, w, 30, oracle.F64, probes=w.probes)
    _, st, series = b200_run(b200, w, 30, probes=w.probes)
    assert rel_l2(st.read_field(b200.FieldComponent.E), ref64.field(oracle.E)) <= TOL_VS_F64
    assert rel_l2(st.read_field(b200.FieldComponent.H), ref64.field(oracle.H)) <= TOL_VS_F64
    assert rel_l2(series, s64) <= TOL_VS_F64


@pytest.mark.parametrize("fused", [False, True])
def test_per_step_protocol_equals_the_batched_run(b200, fused):
    """set_sources -> step(1) -> read_probes every step (runner.rs:522-526; the forcing row travels through the pinned ring
    without a synchronize, the probe row comes back through the pinned mirror) == set_source_table + step(n), bit for bit;
    a changed cell set, a step without forcing and a late re-read of an old tick take the synchronous paths."""
    w = _random_workload(b200, (128, 16, 12), seed=5)
    probes = [(64, 8, 6), (0, 0, 0), (127, 15, 11), (65, 8, 6)]
    n = 24
    _, ref, ref_series = b200_run(b200, w, n, exact=True, probes=probes, fused=fused)
    inst, st = b200_state(b200, w, exact=True, fused=fused)
    assert st.uses_fused == fused
    st.register_probes(probes)
    table = w.source_table(n)
    pts = w.source_points()
    rows = []
    for s in range(n):
        up = inst.begin_update(st)
        for k, p in enumerate(pts):
            up.set_forcing(p, b200.SourceValues(tuple(table[s, k, :3]), tuple(table[s, k, 3:])))
        up.finish()
        rows.append(st.read_probes(s, 1)[0])
    assert np.array_equal(np.stack(rows), ref_series)
    assert np.array_equal(st.read_probes(n - 20, 20), ref_series[n - 20:])       # older ticks: from the device ring
    assert np.array_equal(st.read_field(b200.FieldComponent.E), ref.read_field(b200.FieldComponent.E))
    assert np.array_equal(st.read_field(b200.FieldComponent.H), ref.read_field(b200.FieldComponent.H))
    # another cell set, then none, then the first one again, then both
    other = (3, 2, 1)
    vals = np.array([1.0, 2.0, -3.0, 0.5, 0.0, 0.25], dtype=np.float32)
    for step_pts in ([other], [], pts, pts + [other]):
        up = inst.begin_update(st)
        for p in step_pts:
            up.set_forcing(p, b200.SourceValues(tuple(vals[:3]), tuple(vals[3:])))
        up.finish()
        if step_pts:
            ref.set_source_table(step_pts, np.tile(vals, (1, len(step_pts), 1)))
        ref.step(1)
        assert np.array_equal(st.read_probes(st.tick() - 1, 1), ref.read_probes(ref.tick() - 1, 1))
    assert st.tick() == ref.tick() == n + 4
    assert np.array_equal(st.read_field(b200.FieldComponent.E), ref.read_field(b200.FieldComponent.E))


@pytest.mark.parametrize("fused", [False, True])
def test_probe_rows_read_one_step_behind_equal_the_batched_run(b200, fused):
    """The row of step s read while step s + 1 is already enqueued (fdtd_b200_read_probes waits for the event behind that row's
    pinned copy, not for the stream) == the batched series; rows older than the pinned ring come from the device ring."""
    w = _random_workload(b200, (128, 16, 12), seed=6)
    probes = [(64, 8, 6), (0, 0, 0), (127, 15, 11)]
    n = 30
    _, _, ref_series = b200_run(b200, w, n, exact=True, probes=probes, fused=fused)
    inst, st = b200_state(b200, w, exact=True, fused=fused)
    st.register_probes(probes)
    table = w.source_table(n)
    pts = w.source_points()
    rows = []
    for s in range(n):
        up = inst.begin_update(st)
        for k, p in enumerate(pts):
            up.set_forcing(p, b200.SourceValues(tuple(table[s, k, :3]), tuple(table[s, k, 3:])))
        up.finish()
        if s:
            rows.append(st.read_probes(s - 1, 1)[0])
    rows.append(st.read_probes(n - 1, 1)[0])
    assert np.array_equal(np.stack(rows), ref_series)
    assert np.array_equal(st.read_probes(n - 12, 1)[0], ref_series[n - 12])     # its pinned slot has been reused since


@pytest.mark.parametrize("fused", [False, True])
def test_one_step_and_batched_calls_share_one_probe_series(b200, fused):
    """One-step calls write their probe row into the host-mapped ring, batched calls into the HBM ring: a read that spans both
    kinds of rows, a reset in between and the FDTD_B200_STAGED_IO path all return the batched run's series."""
    w = _random_workload(b200, (128, 16, 12), seed=8)
    probes = [(64, 8, 6), (127, 15, 11), (1, 0, 3)]
    n = 14
    _, _, ref_series = b200_run(b200, w, n, exact=True, probes=probes, fused=fused)
    inst, st = b200_state(b200, w, exact=True, fused=fused)
    st.register_probes(probes)
    for rnd in range(2):
        st.set_source_table(w.source_points(), w.source_table(n))
        for k in (1, 1, 1, 5, 1, 4, 1):
            st.step(k)
        assert st.tick() == n
        assert np.array_equal(st.read_probes(0, n), ref_series)
        assert np.array_equal(st.read_probes(2, 1)[0], ref_series[2]) and np.array_equal(st.read_probes(5, 3), ref_series[5:8])
        st.reset()
        st.register_probes(probes) if rnd == 0 else None
