"""CPML on the device (parity unpinned against the reference, which never runs the recursion): the ψ passes
against the oracle's in-formula ψ (tolerance: the passes add `coeff·psi` after the update, one rounding later than the
oracle's `coeff·(curl − src + psi)`), fused vs two-kernel path bit for bit, z-slabs, probes inside the layers, reset,
checkpoints, and the absorption itself at −40 dB."""
import math

import numpy as np
import pytest

from helpers import b200_run, b200_state, oracle_run, random_fields, rel_l2
from test_parity_gpu import _fields, _random_workload

pytestmark = pytest.mark.gpu

TOL_PML_VS_F64 = 1e-5   # north-star tolerance; measured ≈ 2e-7 on these cases


def _with_pml(b200, w, thickness, staggered=True, axes=(0, 1, 2), sigma_max=2.5):
    w.pml = b200.PmlProfiles.graded(w.solver_config(), thickness, b200.GradedPml(sigma_max=sigma_max), axes=axes, staggered=staggered)
    return w


@pytest.mark.parametrize("size,generic,thickness", [
    ((32, 16, 12), False, 4), ((32, 16, 12), True, 4), ((132, 9, 10), False, 3), ((9, 10, 11), False, 3), ((24, 1, 16), False, 5),
])
def test_pml_passes_match_the_oracle(b200, oracle, size, generic, thickness):
    w = _with_pml(b200, _random_workload(b200, size, seed=sum(size), dx=(1.0, 0.5, 2.0), dt=0.2, n_mat=3), thickness)
    init = random_fields(w, seed=1)
    probes = [(0, 0, 0), (1, min(2, size[1] - 1), 1), (size[0] // 2, size[1] // 2, size[2] // 2), (size[0] - 1, size[1] - 1, size[2] - 1)]
    steps = 30
    ref, sref = oracle_run(oracle, w, steps, oracle.F64, init=init, probes=probes)
    for exact in (True, False):
        _, st, series = b200_run(b200, w, steps, exact=exact, generic=generic, init=init, probes=probes)
        assert st.has_pml()
        e, h = _fields(st, b200)
        assert rel_l2(e, ref.field(oracle.E)) <= TOL_PML_VS_F64 and rel_l2(h, ref.field(oracle.H)) <= TOL_PML_VS_F64
        assert rel_l2(series, sref) <= TOL_PML_VS_F64          # probes inside the layers see the corrected fields
    # and the PML does something: the same run without it differs
    w.pml = None
    _, st0, _ = b200_run(b200, w, steps, exact=True, generic=generic, init=init)
    assert rel_l2(st0.read_field(b200.FieldComponent.E), ref.field(oracle.E)) > 1e-3


def test_pml_u16_index_and_single_axis(b200, oracle):
    from fdtd_b200 import configs

    rng = np.random.default_rng(5)
    size = (16, 12, 10)
    pal = [b200.Material(relative_permittivity=1.0 + 0.01 * k) for k in range(300)]
    ids = rng.integers(0, 300, size[0] * size[1] * size[2]).astype(np.uint32)
    ids[:300] = np.arange(300)
    w = configs.Workload("u16", size, b200.Resolution((1.0, 1.0, 1.0), 0.25), b200.PhysicalConstants.REDUCED, pal, ids,
                         [((8, 6, 5), b200.GaussianPulse(2.0, 1.0).with_amplitudes((0.0, 0.0, 5.0), (1.0, 0.0, 0.0)))], [], 0)
    _with_pml(b200, w, 3, axes=(1,))
    assert w.pml.axes[0] is None and w.pml.axes[2] is None
    ref, _ = oracle_run(oracle, w, 25, oracle.F64)
    _, st, _ = b200_run(b200, w, 25, exact=True)
    assert st.index_bytes == 2
    assert rel_l2(st.read_field(b200.FieldComponent.E), ref.field(oracle.E)) <= TOL_PML_VS_F64


@pytest.mark.parametrize("exact", [True, False])
def test_fused_step_with_pml_equals_two_kernel_path(b200, exact):
    """Fused kernel + (H pass, E recomputed in the touched slabs, E pass) on the ping-pong sets == H kernel, H pass,
    E kernel, E pass in place — bit for bit, in both arithmetic modes, with a source and probes inside the layers."""
    size = (128, 18, 14)
    w = _with_pml(b200, _random_workload(b200, size, seed=3, dx=(1.0, 1.0, 1.0), dt=0.25), 4)
    w.sources.append(((1, 2, 1), b200.ContinousWave(0.1, 0.05).with_amplitudes((1.0, 2.0, -1.0), (0.5, 0.0, 0.25))))   # inside the PML
    init = random_fields(w, seed=8)
    probes = [(0, 0, 0), (2, 3, 1), (64, 9, 7), (127, 17, 13), (126, 1, 12)]
    steps = 11                                           # odd: the ping-pong ends on the other set
    _, a, sa = b200_run(b200, w, steps, exact=exact, init=init, probes=probes)
    _, f, sf = b200_run(b200, w, steps, exact=exact, init=init, probes=probes, fused=True)
    assert f.uses_fused and not a.uses_fused and f.has_pml()
    for comp in (b200.FieldComponent.E, b200.FieldComponent.H):
        assert np.array_equal(a.read_field(comp), f.read_field(comp))
    assert np.array_equal(sa, sf)
    assert np.array_equal(a.read_pml_state(), f.read_pml_state())


@pytest.mark.parametrize("size,thickness,exact,u16", [
    ((128, 18, 14), 4, True, False), ((256, 30, 20), 5, True, False), ((256, 30, 20), 9, False, False),
    ((384, 40, 70), 16, False, False), ((144, 23, 9), 3, True, True),
    # x slabs that cross a 128-cell tile boundary (the recomputed column right of a tile lies in the slab): aligned and padded rows
    ((272, 18, 14), 16, True, False), ((400, 12, 20), 20, False, False), ((520, 10, 9), 16, True, False), ((135, 9, 8), 9, True, False),
])
def test_pml_inside_the_fused_kernel_equals_the_passes(b200, monkeypatch, size, thickness, exact, u16):
    """The fused kernel's PML instantiation (ψ_H ping-pong, ψ_E in place, x-slab ψ staged by cp.async) == the fused kernel
    followed by the slab passes (FDTD_B200_PML_PASSES=1) — bit for bit: fields, probes inside the layers and the ψ state.
    Lattices with several tiles along x and y, several z chunks, layers that are not quad-aligned, both index widths."""
    # exact arithmetic with dx = (1, 0.5, 2) is the multiply mode, with dx = (1, 0.3, 2) the IEEE-division mode
    dx = (1.0, 0.3, 2.0) if (exact and size[0] == 256) else (1.0, 0.5, 2.0)
    w = _with_pml(b200, _random_workload(b200, size, seed=sum(size), dx=dx, dt=0.1), thickness)
    if u16:
        rng = np.random.default_rng(17)
        w.palette = [b200.Material(relative_permittivity=1.0 + 0.01 * m, eletrical_conductivity=0.001 * m) for m in range(300)]
        w.ids = rng.integers(0, 300, w.n_cells).astype(np.uint32)
        w.ids[:300] = np.arange(300)
    w.sources.append(((1, 2, 1), b200.ContinousWave(0.1, 0.05).with_amplitudes((1.0, 2.0, -1.0), (0.5, 0.0, 0.25))))
    w.sources.append(((size[0] - 1, size[1] - 2, size[2] - 1), b200.GaussianPulse(1.0, 0.5).with_amplitudes((0.0, 1.0, 0.0), (0.0, 0.0, 2.0))))
    init = random_fields(w, seed=2)
    probes = [(0, 0, 0), (2, 3, 1), (size[0] // 2, size[1] // 2, size[2] // 2), (size[0] - 1, size[1] - 1, size[2] - 1), (127, 7, 2)]
    steps = 7
    monkeypatch.setenv("FDTD_B200_PML_PASSES", "1")
    _, p, sp = b200_run(b200, w, steps, exact=exact, init=init, probes=probes, fused=True)
    monkeypatch.delenv("FDTD_B200_PML_PASSES")
    _, k, sk = b200_run(b200, w, steps, exact=exact, init=init, probes=probes, fused=True)
    assert p.uses_fused and k.uses_fused and k.pml_in_kernel() and not p.pml_in_kernel()
    assert k.index_bytes == (2 if u16 else 1)
    for comp in (b200.FieldComponent.E, b200.FieldComponent.H):
        assert np.array_equal(p.read_field(comp), k.read_field(comp))
    assert np.array_equal(sp, sk)
    assert np.array_equal(p.read_pml_state(), k.read_pml_state())
    assert k.kernel_launches < p.kernel_launches


def test_pml_layers_the_fused_tiling_cannot_hold_fall_back_to_the_passes(b200):
    """More than 32 PML columns in one 128-cell tile: the handle runs the slab passes after the fused kernel instead."""
    size = (128, 12, 10)
    w = _with_pml(b200, _random_workload(b200, size, seed=9), 20, axes=(0,))
    init = random_fields(w, seed=3)
    _, a, _ = b200_run(b200, w, 5, init=init)
    _, f, _ = b200_run(b200, w, 5, init=init, fused=True)
    assert f.uses_fused and f.has_pml() and not f.pml_in_kernel()
    for comp in (b200.FieldComponent.E, b200.FieldComponent.H):
        assert np.array_equal(a.read_field(comp), f.read_field(comp))


@pytest.mark.parametrize("size,cuts,thickness", [
    ((128, 16, 24), (9,), 5),            # the cut lies outside the z layers
    ((128, 16, 24), (3, 21), 5),         # both cuts run through a z layer: the plane above a slab is a PML plane
    ((256, 10, 12), (1, 3, 10), 4),      # one- and two-plane slabs (no interior launch), four slabs
])
def test_pml_on_fused_slabs_over_the_peer_memory_transport_equals_one_domain(b200, size, cuts, thickness):
    """Fused z-slabs with the CPML inside the kernel (fdtd_b200_p2p_*; all slab handles in one process on one GPU): every
    slab recomputes H' of the plane above it and needs that plane's ψ_H, which the upper slab pushes with its boundary
    planes.  Bit-identical to the whole-domain fused run, including sources and probes in the layers, and after a reset."""
    w = _with_pml(b200, _random_workload(b200, size, seed=sum(size) + len(cuts), dx=(1.0, 1.0, 1.0), dt=0.25), thickness)
    nz = size[2]
    w.sources.append(((2, 1, cuts[0]), b200.GaussianPulse(0.2, 0.5).with_amplitudes((1, 2, 3), (-1, 0.5, 2))))
    probes = [(1, 1, cuts[0] - 1), (3, 2, cuts[0]), (size[0] - 1, size[1] - 1, nz - 1), (size[0] // 2, size[1] // 2, nz // 2)]
    init = random_fields(w, seed=6)
    steps = 7
    tab = w.source_table(steps)
    bounds = list(zip((0,) + tuple(cuts), tuple(cuts) + (nz,)))
    _, whole, s_whole = b200_run(b200, w, steps, exact=True, init=init, probes=probes, fused=True)
    assert whole.pml_in_kernel()
    slabs = []
    for a, b in bounds:
        st = b200.FdtdB200Backend(exact_arithmetic=True, fused_step=True, z_slab=(a, b)).create_instance(w.solver_config(), w.domain()).create_state()
        assert st.uses_fused and st.has_pml() and st.pml_in_kernel()
        slabs.append(st)
    blobs = [st.p2p_export() for st in slabs]
    for k, st in enumerate(slabs):
        st.p2p_connect(blobs[k - 1] if k > 0 else None, blobs[k + 1] if k + 1 < len(slabs) else None)
    for rnd in range(2):
        for st, (a, b) in zip(slabs, bounds):
            if rnd:
                st.reset()
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
            assert np.array_equal(got, whole.read_field(comp)), (rnd, comp)
        assert np.array_equal(sum(st.read_probes(0, steps) for st in slabs), s_whole)


def test_pml_on_two_slabs_equals_one_domain(b200):
    from fdtd_b200 import _lib as L

    size, cut = (32, 12, 16), 3                          # the cut runs through the low-z layer
    w = _with_pml(b200, _random_workload(b200, size, seed=11), 5)
    init = random_fields(w, seed=4)
    steps = 9
    _, whole, _ = b200_run(b200, w, steps, init=init)
    nz = size[2]
    _, lo = b200_state(b200, w, z_slab=(0, cut))
    _, hi = b200_state(b200, w, z_slab=(cut, nz))
    assert lo.has_pml() and hi.has_pml()
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
        lo.update_h(0, cut); hi.update_h(0, 1); hi.update_h(1, nz - cut)          # plane ranges, as the overlapping stepper issues them
        h_down()
        lo.update_e(0, cut - 1); lo.update_e(cut - 1, cut); hi.update_e(0, nz - cut)
        e_up()
        lo.end_step(); hi.end_step()
    for comp in (b200.FieldComponent.E, b200.FieldComponent.H):
        got = np.concatenate([lo.read_field(comp), hi.read_field(comp)], axis=0)
        assert np.array_equal(got, whole.read_field(comp))


def test_pml_state_reset_checkpoint_and_errors(b200, tmp_path):
    from fdtd_b200 import _lib as L
    from fdtd_b200 import io as fio

    size = (128, 12, 10)
    w = _with_pml(b200, _random_workload(b200, size, seed=9, dx=(1.0, 1.0, 1.0), dt=0.25), 3)
    tab = w.source_table(20)

    def fresh(fused):
        return b200.FdtdB200Backend(exact_arithmetic=True, fused_step=fused).create_instance(w.solver_config(), w.domain()).create_state()

    for fused in (False, True):
        ref = fresh(fused)
        ref.set_source_table(w.source_points(), tab)
        ref.step(20)
        a = fresh(fused)
        n_psi = a.read_pml_state().size
        assert n_psi == 4 * (6 * 12 * 10 + 128 * 6 * 10 + 128 * 12 * 6) + 4 * (12 * 10 + 128 * 10 + 128 * 12)   # E slabs 3+3, H one more on the low side
        assert not a.read_pml_state().any()
        a.set_source_table(w.source_points(), tab)
        a.step(7)
        assert a.read_pml_state().any()
        path = str(tmp_path / f"ck{int(fused)}.bin")
        fio.save_checkpoint(a, path)
        b = fresh(fused)
        fio.load_checkpoint(b, path)
        b.set_source_table(w.source_points(), tab[7:])
        b.step(13)
        for comp in (b200.FieldComponent.E, b200.FieldComponent.H):
            assert np.array_equal(b.read_field(comp), ref.read_field(comp))
        assert np.array_equal(b.read_pml_state(), ref.read_pml_state())
        a.reset()                                          # create_state semantics: ψ back to zero
        assert not a.read_pml_state().any() and a.tick() == 0
        a.set_source_table(w.source_points(), tab)
        a.step(20)
        assert np.array_equal(a.read_field(b200.FieldComponent.E), ref.read_field(b200.FieldComponent.E))
        a.set_pml(None)                                    # cleared
        assert not a.has_pml() and a.read_pml_state().size == 0
        w2 = _random_workload(b200, size, seed=9, dx=(1.0, 1.0, 1.0), dt=0.25)
        plain = b200.FdtdB200Backend(exact_arithmetic=True, fused_step=fused).create_instance(w2.solver_config(), w2.domain()).create_state()
        with pytest.raises(ValueError):
            fio.load_checkpoint(plain, path)               # ψ count mismatch

    st = fresh(False)
    dp = np.zeros(5)
    with pytest.raises(L.Fdtd_b200Error):
        L.check(st._lib.fdtd_b200_set_pml(st._h, 3, None, None, None, None), st._h)
    nan = np.full(128, np.nan)
    ptr = nan.ctypes.data_as(L._dp)
    with pytest.raises(L.Fdtd_b200Error):
        L.check(st._lib.fdtd_b200_set_pml(st._h, 0, ptr, ptr, None, None), st._h)
    # fused slab handles carry the PML inside the kernel only; ψ_H of the plane above the slab arrives over the peer-memory
    # transport, so stepping the slab's last plane without it is refused, and so are layers the kernel's tiling cannot hold
    slab = b200.FdtdB200Backend(fused_step=True, z_slab=(0, 5)).create_instance(w2.solver_config(), w2.domain()).create_state()
    assert slab.uses_fused
    slab.set_pml(w.pml)
    assert slab.has_pml() and slab.pml_in_kernel()
    with pytest.raises(L.Fdtd_b200Error):
        slab.fused_update(0, 5)
    with pytest.raises(L.Fdtd_b200Error):
        slab.set_pml(b200.PmlProfiles.graded(w.solver_config(), 20, b200.GradedPml(), axes=(0,)))
    del dp


def _pulse(b200, n, thickness, steps, probes_rel, fused, pml=True):
    from fdtd_b200 import configs

    c = n // 2
    src = b200.GaussianPulse(12.5, 2.5).with_amplitudes((0.0, 0.0, 50.0), (0.0, 0.0, 0.0))
    w = configs.Workload("pulse", (n, n, n), b200.Resolution((1.0, 1.0, 1.0), 0.25), b200.PhysicalConstants.REDUCED,
                         [b200.Material.VACUUM], None, [((c, c, c), src)], [], 0)
    if pml:
        _with_pml(b200, w, thickness)
    probes = [(c + dx, c + dy, c + dz) for (dx, dy, dz) in probes_rel]
    _, st, series = b200_run(b200, w, steps, probes=probes, fused=fused)
    return st, series


@pytest.mark.parametrize("fused", [False, True])
def test_cpml_absorbs_on_the_device(b200, fused):
    """128³ lattice with 10-cell layers against a 384³ lattice whose walls the pulse cannot reach and return from within
    the run: reflection below −40 dB.  Same measurement as tests/test_pml.py::test_oracle_cpml_absorbs."""
    probes = [(40, 0, 0), (0, 40, 0), (0, 0, 40), (30, 30, 0), (25, 25, 25), (45, 12, 33)]
    steps = 900                                            # the pulse crosses the small box ≈ 1.8 times
    _, big = _pulse(b200, 384, 0, steps, probes, fused, pml=False)
    st, small = _pulse(b200, 128, 10, steps, probes, fused)
    assert st.uses_fused == fused and st.has_pml()
    ref = np.abs(big[..., :3]).max()
    refl = 20 * math.log10(np.abs(small[..., :3] - big[..., :3]).max() / ref)
    assert refl < -40.0, refl
    _, wall = _pulse(b200, 128, 10, steps, probes, fused, pml=False)
    assert 20 * math.log10(np.abs(wall[..., :3] - big[..., :3]).max() / ref) > -25.0
