"""CUDA scene rasteriser (kernels_scene.cuh) against oracle/scene_ref.py — integer material indexing, so bit-exact —
and BASELINE configs[4] (NEC dipole) end to end: deck → cylinders → device-rasterised index lattice → fields and probe
series against the CPU oracle."""
import dataclasses

import numpy as np
import pytest

from helpers import oracle_run, rel_l2

pytestmark = pytest.mark.gpu


def _mixed_scene(lattice, dx=(0.5, 0.25, 0.75), mins=(-3.0, -2.0, -4.0), rot=(0.9, 0.1, -0.3, 0.2)):
    from fdtd_b200.scene import CoordinateTransformations, SceneDomain, Shape
    from fdtd_b200.solver import Material, Resolution

    nx, ny, nz = lattice
    tf = CoordinateTransformations.for_fdtd(Resolution(dx, 0.1), lattice, rotation=rot, aabb_mins=mins)
    centre = tf.transform_point_from_solver_to_world((nx // 2, ny // 2, nz // 2)).astype(np.float64)
    ext = min(nx * dx[0], ny * dx[1], nz * dx[2])
    rng = np.random.default_rng(7)
    mats = [Material(relative_permittivity=2.0 + k) for k in range(5)] + [Material(eletrical_conductivity=10.0)]
    shapes = []
    for k in range(40):
        c = tuple(centre + rng.uniform(-0.5, 0.5, 3) * ext)
        q = tuple(rng.standard_normal(4))
        m = mats[k % len(mats)]
        kind = k % 3
        if kind == 0:
            shapes.append(Shape.ball(rng.uniform(0.05, 0.2) * ext, m, c))
        elif kind == 1:
            shapes.append(Shape.cuboid(rng.uniform(0.03, 0.2, 3) * ext, m, c, q))
        else:
            shapes.append(Shape.cylinder(rng.uniform(0.05, 0.3) * ext, rng.uniform(0.02, 0.1) * ext, m, c, q))
    return SceneDomain(shapes, tf)


@pytest.mark.parametrize("lattice", [(64, 40, 24), (37, 21, 13), (5, 1, 3), (1, 1, 1)])
def test_rasteriser_is_bit_identical_to_the_oracle(b200, lattice):
    from oracle import scene_ref

    scene = _mixed_scene(lattice)
    ref = scene_ref.rasterize_scene_domain(scene)
    got = scene.rasterize()
    assert got.dtype == np.uint32 and np.array_equal(got, ref)
    if lattice[2] > 4:
        assert len(np.unique(ref)) > 3                                  # the scene actually exercises several labels
        assert np.array_equal(scene.rasterize(z_begin=2, z_end=lattice[2] - 1), scene_ref.rasterize_scene_domain(scene, 2, lattice[2] - 1))


def test_empty_scene_and_argument_errors(b200):
    import ctypes as C

    from fdtd_b200 import _lib as L
    from fdtd_b200.scene import CoordinateTransformations, SceneDomain, Shape
    from fdtd_b200.solver import Material, Resolution

    tf = CoordinateTransformations.for_fdtd(Resolution((1.0, 1.0, 1.0), 0.25), (8, 4, 2))
    assert not SceneDomain([], tf).rasterize().any()
    scene = SceneDomain([Shape.ball(1.0, Material(relative_permittivity=2.0), (1.0, 1.0, 1.0))], tf)
    with pytest.raises(L.Fdtd_b200Error):
        scene.rasterize(z_begin=2, z_end=2)
    bad = scene.c_shapes()
    bad[0].kind = 9
    out = np.zeros(64, dtype=np.uint32)
    rc = L.lib().fdtd_b200_rasterize_scene(0, (C.c_uint64 * 3)(8, 4, 2), 0, 2, scene.c_transform(), bad, 1, 0, out.ctypes.data_as(C.POINTER(C.c_uint32)))
    assert rc == L.EINVAL


@pytest.mark.parametrize("fused", [False, True])
def test_scene_domain_builds_the_device_index_without_a_host_lattice(b200, fused):
    from oracle import scene_ref

    lattice = (128, 24, 20)
    scene = _mixed_scene(lattice, dx=(1.0, 1.0, 1.0), mins=(0.0, 0.0, 0.0), rot=(1.0, 0.0, 0.0, 0.0))
    from fdtd_b200.solver import FdtdSolverConfig, PhysicalConstants, Resolution

    cfg = FdtdSolverConfig.from_lattice(lattice, Resolution((1.0, 1.0, 1.0), 0.25), PhysicalConstants.REDUCED)
    ref = scene_ref.rasterize_scene_domain(scene)
    st = b200.FdtdB200Backend(fused_step=fused).create_instance(cfg, scene).create_state()
    assert np.array_equal(st.material_ids(), ref)
    st.close()
    # z-slab handle: its own planes (the fused slab also fills the plane above, checked through a run in test_fused_gpu)
    st = b200.FdtdB200Backend(fused_step=fused, z_slab=(5, 14)).create_instance(cfg, scene).create_state()
    assert np.array_equal(st.material_ids(), ref.reshape(20, -1)[5:14].reshape(-1))
    st.close()


def test_config5_nec_dipole_matches_the_oracle(b200, oracle):
    from fdtd_b200 import configs
    from oracle import scene_ref

    steps = 300
    w = configs.config5_nec_dipole(96, steps)
    ids = scene_ref.rasterize_scene_domain(w.scene)
    w_ref = dataclasses.replace(w, ids=ids, scene=None)
    ref64, s64 = oracle_run(oracle, w_ref, steps, oracle.F64, probes=w.probes)
    ref32, s32 = oracle_run(oracle, w_ref, steps, oracle.F32, probes=w.probes)
    for exact in (False, True):
        inst = b200.FdtdB200Backend(exact_arithmetic=exact).create_instance(w.solver_config(), w.domain())
        st = inst.create_state()
        assert np.array_equal(st.material_ids(), ids)                     # deck → cylinders → device index: bit-exact
        st.register_probes(w.probes)
        st.set_source_table(w.source_points(), w.source_table(steps))
        st.step(steps)
        e, h = st.read_field(b200.FieldComponent.E), st.read_field(b200.FieldComponent.H)
        series = st.read_probes(0, steps)
        if exact:   # dx is not a power of two → true division mode; still bit-identical to the fp32 oracle
            assert np.array_equal(e, ref32.field(oracle.E).astype(np.float32)) and np.array_equal(h, ref32.field(oracle.H).astype(np.float32))
            assert np.array_equal(series, s32.astype(np.float32))
        assert rel_l2(e, ref64.field(oracle.E)) <= 1e-5 and rel_l2(h, ref64.field(oracle.H)) <= 1e-5
        assert rel_l2(series, s64) <= 1e-5
        assert np.abs(series).max() > 0                                   # the pulse reached the probe ring
        st.close()


def test_config5_full_size_wire_and_short_run(b200, oracle):
    """256³ as BASELINE names it: the index lattice against the oracle rasteriser on the planes that hold the wire, the
    whole-lattice label histogram, and 28 steps against the f64 oracle on a sub-box around the feed."""
    from fdtd_b200 import configs
    from oracle import scene_ref

    steps = 28
    w = configs.config5_nec_dipole(256, steps)
    st = b200.FdtdB200Backend(fused_step=True).create_instance(w.solver_config(), w.domain()).create_state()
    ids = st.material_ids().reshape(256, 256, 256)
    assert np.array_equal(ids[120:136].reshape(-1), scene_ref.rasterize_scene_domain(w.scene, 120, 136))
    assert np.bincount(ids.reshape(-1)).tolist() == [256 ** 3 - 60, 60]
    assert not ids[:120].any() and not ids[136:].any()
    st.set_source_table(w.source_points(), w.source_table(steps))
    st.step(steps)
    assert st.uses_fused and st.tick() == steps
    # the stencil's domain of dependence grows one cell per step: after 28 steps everything non-zero lies within 28 cells of
    # the feed cell, strictly inside a 64-cell-wide column around the wire (whose own Neumann faces are then never felt)
    sub = dataclasses.replace(w, lattice=(64, 256, 64), scene=None,
                              ids=np.ascontiguousarray(ids[96:160, :, 96:160]).reshape(-1),
                              sources=[((32, 128, 32), w.sources[0][1])])
    ref64, _ = oracle_run(oracle, sub, steps, oracle.F64)
    e = st.read_field(b200.FieldComponent.E, (96, 0, 96), (160, 256, 160))
    assert rel_l2(e, ref64.field(oracle.E)) <= 1e-5
    full_e = st.read_field(b200.FieldComponent.E)
    outside = full_e.copy()
    outside[96:160, :, 96:160] = 0
    assert not outside.any()
    st.close()
