"""GPU tests of row padding (include/fdtd_b200.h, fdtd_b200_row_pitch): lattices whose nx is not a multiple of 16 — the usual
case, the reference sizes its lattice as ceil(size/dx) (fdtd/mod.rs:24-29) — are stored with padded rows so that the float4 /
TMA / fused kernels run on them.  The pad cells must be invisible: every result is bit-identical to the fp32 oracle (exact
arithmetic) and to the unpadded handle (FDTD_B200_NO_ROW_PADDING=1, the one-thread-per-cell kernels), and everything that
crosses the ABI stays in terms of the caller's nx × ny × nz lattice."""
import ctypes as C

import numpy as np
import pytest

from helpers import b200_run, b200_state, oracle_run, random_fields, rel_l2
from test_parity_gpu import _random_workload
from test_pml_gpu import _with_pml

pytestmark = pytest.mark.gpu


def _pitch(nx):
    if nx % 16 == 0 or nx < 64:
        return nx
    return 128 if 96 < nx < 128 else (nx + 15) // 16 * 16


def _edge_workload(b200, size, seed, **kw):
    """Random materials, sources and probes on the last lattice column (the cells whose x+1 neighbour is a pad cell)."""
    w = _random_workload(b200, size, seed=seed, **kw)
    nx, ny, nz = size
    w.sources.append(((nx - 1, ny // 2, nz // 2), b200.GaussianPulse(0.3, 0.4).with_amplitudes((0.5, 1.5, -2.0), (3.0, -1.0, 0.25))))
    w.sources.append(((nx - 2, 0, nz - 1), b200.ContinousWave(0.0, 0.11).with_amplitudes((1.0, 0.0, 2.0), (0.0, 2.0, 0.0))))
    probes = [(nx - 1, ny - 1, nz - 1), (nx - 1, 0, 0), (nx - 2, ny // 2, nz // 2), (0, 0, 0), (nx // 2, ny // 2, nz // 2), (min(127, nx - 1), min(7, ny - 1), 0)]
    return w, probes


@pytest.mark.parametrize("size,fused", [
    ((130, 9, 7), True), ((200, 12, 6), True), ((250, 17, 9), True), ((100, 10, 5), True), ((129, 3, 3), True), ((333, 8, 20), True),
    ((257, 20, 5), True), ((127, 8, 70), True),
    ((65, 5, 4), False), ((200, 12, 6), False), ((130, 9, 7), False), ((90, 6, 5), True),
])
def test_unaligned_rows_are_bit_identical_to_the_f32_oracle(b200, oracle, size, fused):
    w, probes = _edge_workload(b200, size, seed=sum(size) + 1)
    init = random_fields(w, seed=10)
    steps = 5
    ref, s32 = oracle_run(oracle, w, steps, oracle.F32, init=init, probes=probes)
    _, st, series = b200_run(b200, w, steps, exact=True, init=init, probes=probes, fused=fused)
    assert st.row_pitch == _pitch(size[0]) and st.row_pitch != size[0]
    assert st.uses_vec4 and st.uses_fused == (fused and st.row_pitch >= 128)
    assert np.array_equal(st.read_field(b200.FieldComponent.E), ref.field(oracle.E).astype(np.float32))
    assert np.array_equal(st.read_field(b200.FieldComponent.H), ref.field(oracle.H).astype(np.float32))
    assert np.array_equal(series, s32.astype(np.float32))
    # the caller's view of the material lattice and of the table is unchanged by the pad row
    assert np.array_equal(st.material_ids(), w.ids)
    assert st.coefficient_table().shape[0] == len(w.palette)
    # boxes are clipped at the caller's nx
    box = st.read_field(b200.FieldComponent.E, (size[0] - 3, 0, 0), (size[0] + 40, 2, 1))
    assert box.shape == (1, min(2, size[1]), 3, 3)
    assert np.array_equal(box, ref.field(oracle.E).astype(np.float32)[:1, :2, size[0] - 3:])


def test_fast_mode_on_padded_rows_meets_the_tolerance(b200, oracle):
    size = (200, 24, 20)
    w, probes = _edge_workload(b200, size, seed=5, dx=(1.0, 1.0, 1.0), dt=0.25)
    steps = 60
    ref, s64 = oracle_run(oracle, w, steps, oracle.F64, probes=probes)
    _, st, series = b200_run(b200, w, steps, probes=probes, fused=True)
    assert st.uses_fused and st.row_pitch == 208
    assert rel_l2(st.read_field(b200.FieldComponent.E), ref.field(oracle.E)) <= 1e-5
    assert rel_l2(st.read_field(b200.FieldComponent.H), ref.field(oracle.H)) <= 1e-5
    assert rel_l2(series, s64) <= 1e-5


def test_padding_is_off_with_the_environment_switch_and_equal_to_it(b200, monkeypatch):
    size = (150, 10, 8)
    w, probes = _edge_workload(b200, size, seed=3)
    init = random_fields(w, seed=1)
    _, a, sa = b200_run(b200, w, 6, exact=True, init=init, probes=probes, fused=True)
    monkeypatch.setenv("FDTD_B200_NO_ROW_PADDING", "1")
    _, b, sb = b200_run(b200, w, 6, exact=True, init=init, probes=probes, fused=True)
    monkeypatch.delenv("FDTD_B200_NO_ROW_PADDING")
    assert a.row_pitch == 160 and a.uses_fused and b.row_pitch == 150 and not b.uses_fused and not b.uses_vec4
    for comp in (b200.FieldComponent.E, b200.FieldComponent.H):
        assert np.array_equal(a.read_field(comp), b.read_field(comp))
    assert np.array_equal(sa, sb)
    be = b200.FdtdB200Backend(fused_step=True)
    assert be.memory_required(w.solver_config()) > 12 * 4 * 160 * 10 * 8       # the pitch, two buffer sets


@pytest.mark.parametrize("size,thickness,exact,u16", [((150, 18, 14), 4, True, False), ((200, 30, 20), 9, False, False), ((270, 23, 9), 3, True, True)])
def test_pml_on_padded_rows_equals_the_unpadded_handle(b200, monkeypatch, size, thickness, exact, u16):
    """CPML inside the fused kernel's padded instantiation (x slab ending at the caller's nx - 1, whose x+1 neighbour is a pad
    cell) == the slab passes after the fused kernel == the unpadded handle's one-thread-per-cell kernels + passes."""
    w, probes = _edge_workload(b200, size, seed=sum(size), dx=(1.0, 0.5, 2.0), dt=0.1)
    w = _with_pml(b200, w, thickness)
    if u16:
        rng = np.random.default_rng(17)
        w.palette = [b200.Material(relative_permittivity=1.0 + 0.01 * m, eletrical_conductivity=0.001 * m) for m in range(300)]
        w.ids = rng.integers(0, 300, w.n_cells).astype(np.uint32)
        w.ids[:300] = np.arange(300)
    init = random_fields(w, seed=2)
    steps = 7
    _, k, sk = b200_run(b200, w, steps, exact=exact, init=init, probes=probes, fused=True)
    monkeypatch.setenv("FDTD_B200_PML_PASSES", "1")
    _, p, sp = b200_run(b200, w, steps, exact=exact, init=init, probes=probes, fused=True)
    monkeypatch.delenv("FDTD_B200_PML_PASSES")
    assert k.uses_fused and k.pml_in_kernel() and p.uses_fused and not p.pml_in_kernel() and k.row_pitch != size[0]
    for comp in (b200.FieldComponent.E, b200.FieldComponent.H):
        assert np.array_equal(k.read_field(comp), p.read_field(comp))
    assert np.array_equal(sk, sp)
    assert np.array_equal(k.read_pml_state(), p.read_pml_state())
    if exact:
        monkeypatch.setenv("FDTD_B200_NO_ROW_PADDING", "1")
        _, u, su = b200_run(b200, w, steps, exact=True, init=init, probes=probes, fused=True)
        monkeypatch.delenv("FDTD_B200_NO_ROW_PADDING")
        assert u.row_pitch == size[0] and not u.uses_fused
        for comp in (b200.FieldComponent.E, b200.FieldComponent.H):
            assert np.array_equal(k.read_field(comp), u.read_field(comp))
        assert np.array_equal(sk, su)


@pytest.mark.parametrize("size,cuts,pml", [((200, 16, 24), (9,), 0), ((150, 10, 12), (1, 3, 10), 0), ((200, 16, 24), (3, 21), 5),
                                           ((130, 16, 24), (3, 21), 5)])   # x slab 125..129 crosses the tile boundary at 128: ghost ψ of the right column
def test_fused_slabs_with_padded_rows_equal_one_domain(b200, size, cuts, pml):
    """z-slabs over the peer-memory transport (all handles in this process): halo planes, pushed index plane and pushed ψ_H carry
    the pitch on both sides; bit-identical to the whole-domain run, which is bit-identical to the oracle above."""
    w, probes = _edge_workload(b200, size, seed=sum(size) + len(cuts), dx=(1.0, 1.0, 1.0), dt=0.25)
    if pml:
        w = _with_pml(b200, w, pml)
    nz = size[2]
    init = random_fields(w, seed=6)
    steps = 7
    tab = w.source_table(steps)
    bounds = list(zip((0,) + tuple(cuts), tuple(cuts) + (nz,)))
    _, whole, s_whole = b200_run(b200, w, steps, exact=True, init=init, probes=probes, fused=True)
    slabs = []
    for a, b in bounds:
        st = b200.FdtdB200Backend(exact_arithmetic=True, fused_step=True, z_slab=(a, b)).create_instance(w.solver_config(), w.domain()).create_state()
        assert st.uses_fused and st.row_pitch == _pitch(size[0])
        slabs.append(st)
    blobs = [st.p2p_export() for st in slabs]
    for k, st in enumerate(slabs):
        st.p2p_connect(blobs[k - 1] if k > 0 else None, blobs[k + 1] if k + 1 < len(slabs) else None)
    for st, (a, b) in zip(slabs, bounds):
        st.write_field(b200.FieldComponent.E, init[0][a:b], (0, 0, a), (size[0], size[1], b))
        st.write_field(b200.FieldComponent.H, init[1][a:b], (0, 0, a), (size[0], size[1], b))
        st.register_probes(probes)
        st.set_source_table(w.source_points(), tab)
    for st in slabs:
        st.p2p_refresh()
    for _ in range(steps):
        for st in slabs:
            st.p2p_step(1)
    for st in slabs:
        st.sync()
    for comp in (b200.FieldComponent.E, b200.FieldComponent.H):
        got = np.concatenate([st.read_field(comp) for st in slabs], axis=0)
        assert np.array_equal(got, whole.read_field(comp)), comp
    assert np.array_equal(sum(st.read_probes(0, steps) for st in slabs), s_whole)


def test_two_kernel_slabs_with_padded_rows_equal_one_domain(b200):
    from fdtd_b200 import _lib as L

    size, cut = (90, 12, 16), 7
    w, _ = _edge_workload(b200, size, seed=11)
    init = random_fields(w, seed=4)
    steps = 6
    _, whole, _ = b200_run(b200, w, steps, init=init)
    nz = size[2]
    _, lo = b200_state(b200, w, z_slab=(0, cut))
    _, hi = b200_state(b200, w, z_slab=(cut, nz))
    assert lo.row_pitch == 96 and lo.halo_plane(L.HALO_SEND_EX_TOP)[1] == 96 * 12
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

    h_down(); e_up()
    for _ in range(steps):
        lo.update_h(0, cut); hi.update_h(0, nz - cut)
        h_down()
        lo.update_e(0, cut); hi.update_e(0, nz - cut)
        e_up()
        lo.end_step(); hi.end_step()
    for comp in (b200.FieldComponent.E, b200.FieldComponent.H):
        assert np.array_equal(np.concatenate([lo.read_field(comp), hi.read_field(comp)], axis=0), whole.read_field(comp))


def test_scene_projection_checkpoint_and_errors_on_padded_rows(b200, tmp_path, monkeypatch):
    from fdtd_b200 import _lib as L
    from fdtd_b200 import io as fio
    from fdtd_b200.solver import FdtdSolverConfig, PhysicalConstants, Resolution
    from oracle import scene_ref
    from test_scene_gpu import _mixed_scene

    # scene → device index lattice, written straight into the pitched rows
    lattice = (150, 24, 20)
    scene = _mixed_scene(lattice, dx=(1.0, 1.0, 1.0), mins=(0.0, 0.0, 0.0), rot=(1.0, 0.0, 0.0, 0.0))
    cfg = FdtdSolverConfig.from_lattice(lattice, Resolution((1.0, 1.0, 1.0), 0.25), PhysicalConstants.REDUCED)
    ref_ids = scene_ref.rasterize_scene_domain(scene)
    st = b200.FdtdB200Backend(fused_step=True).create_instance(cfg, scene).create_state()
    assert st.row_pitch == 160 and np.array_equal(st.material_ids(), ref_ids)
    st.close()
    st = b200.FdtdB200Backend(fused_step=True, z_slab=(5, 14)).create_instance(cfg, scene).create_state()
    assert np.array_equal(st.material_ids(), ref_ids.reshape(20, -1)[5:14].reshape(-1))
    st.close()

    # projection: scaled by the caller's nx, equal to the unpadded handle's image
    size = (150, 40, 6)
    w, probes = _edge_workload(b200, size, seed=8)
    init = random_fields(w, seed=3)
    imgs = []
    for pad in (True, False):
        if not pad:
            monkeypatch.setenv("FDTD_B200_NO_ROW_PADDING", "1")
        inst, s = b200_state(b200, w, exact=True, fused=True)
        monkeypatch.delenv("FDTD_B200_NO_ROW_PADDING", raising=False)
        assert (s.row_pitch != size[0]) == pad
        s.write_field(b200.FieldComponent.E, init[0])
        s.write_field(b200.FieldComponent.H, init[1])
        s.step(2)
        target = b200.ImageTarget(size[0], size[1])
        params = b200.ProjectionParameters(projection=np.diag([1.0, 1.0, 1.0, 1.0]).astype(np.float32),
                                           color_map=(0.25 * np.eye(4)).astype(np.float32) + np.float32(0.1), field=b200.FieldComponent.E)
        proj = inst.create_projection(s, target, params)
        pp = inst.begin_projection_pass(s)
        pp.add_projection(proj)
        pp.finish()
        out = []
        target.with_image_buffer(lambda img: out.append(np.array(img)))
        imgs.append(out[0])
    assert np.array_equal(imgs[0], imgs[1]) and len(np.unique(imgs[0])) > 16

    # checkpoint round trip (odd step count: the fused ping-pong parity differs on resume)
    tab = w.source_table(12)

    def fresh():
        return b200.FdtdB200Backend(exact_arithmetic=True, fused_step=True).create_instance(w.solver_config(), w.domain()).create_state()

    ref = fresh(); ref.set_source_table(w.source_points(), tab); ref.step(12)
    a = fresh(); a.set_source_table(w.source_points(), tab); a.step(5)
    path = str(tmp_path / "ck.bin")
    fio.save_checkpoint(a, path, planes_per_chunk=3)
    b = fresh()
    fio.load_checkpoint(b, path, planes_per_chunk=4)
    b.set_source_table(w.source_points(), tab[5:]); b.step(7)
    for comp in (b200.FieldComponent.E, b200.FieldComponent.H):
        assert np.array_equal(b.read_field(comp), ref.read_field(comp))

    # cell indices are the caller's: the first index past the lattice is rejected although the pitched arrays are longer
    n_cells = size[0] * size[1] * size[2]
    cells = np.array([n_cells], dtype=np.uint64)
    vals = np.zeros(6, dtype=np.float32)
    lib = b200.lib()
    assert lib.fdtd_b200_set_source_table(b._h, 1, cells.ctypes.data_as(C.POINTER(C.c_uint64)), 1, vals.ctypes.data_as(C.POINTER(C.c_float))) == L.EINVAL
    assert lib.fdtd_b200_register_probes(b._h, 1, cells.ctypes.data_as(C.POINTER(C.c_uint64))) == L.EINVAL
    cells[0] = n_cells - 1
    assert lib.fdtd_b200_register_probes(b._h, 1, cells.ctypes.data_as(C.POINTER(C.c_uint64))) == L.OK


def test_a_full_coefficient_table_leaves_no_room_for_the_pad_row(b200, oracle):
    """65 536 materials: the table has no free row for the pad cells, so the handle keeps the caller's nx (no padding) and runs the
    one-thread-per-cell kernels — still bit-identical to the oracle."""
    size = (70, 5, 4)
    w, probes = _edge_workload(b200, size, seed=4)
    rng = np.random.default_rng(3)
    w.palette = [b200.Material(relative_permittivity=1.0 + 1e-4 * m) for m in range(65536)]
    w.ids = rng.integers(0, 65536, w.n_cells).astype(np.uint32)
    w.ids[0] = 0
    init = random_fields(w, seed=5)
    ref, s32 = oracle_run(oracle, w, 3, oracle.F32, init=init, probes=probes)
    _, st, series = b200_run(b200, w, 3, exact=True, init=init, probes=probes, fused=True)
    assert st.row_pitch == 70 and not st.uses_vec4 and st.index_bytes == 2
    assert np.array_equal(st.read_field(b200.FieldComponent.E), ref.field(oracle.E).astype(np.float32))
    assert np.array_equal(series, s32.astype(np.float32))
