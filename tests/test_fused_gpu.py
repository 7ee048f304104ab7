"""GPU tests of the fused single-pass step (csrc/kernels_fused.cuh, FDTD_B200_FUSED_STEP): H and E half-steps in one
kernel on ping-pong buffers must be bit-identical to the fp32 oracle (exact mode) and to the two-kernel path."""
import os

import numpy as np
import pytest

from helpers import oracle_run, random_fields
from test_parity_gpu import _random_workload

pytestmark = pytest.mark.gpu


@pytest.fixture
def tune():
    old = os.environ.get("FDTD_B200_TUNE")

    def set_(v):
        if v is None:
            os.environ.pop("FDTD_B200_TUNE", None)
        else:
            os.environ["FDTD_B200_TUNE"] = v

    yield set_
    set_(old)


def run(b200, w, steps, exact, fused, init=None, probes=None):
    be = b200.FdtdB200Backend(exact_arithmetic=exact, fused_step=fused)
    inst = be.create_instance(w.solver_config(), w.domain())
    st = inst.create_state()
    if init is not None:
        st.write_field(b200.FieldComponent.E, init[0])
        st.write_field(b200.FieldComponent.H, init[1])
    if probes:
        st.register_probes(probes)
    if w.sources:
        st.set_source_table(w.source_points(), w.source_table(steps))
    st.step(steps)
    return st, (st.read_probes(0, steps) if probes else None)


@pytest.mark.parametrize("size,zc", [
    ((128, 8, 6), 0), ((128, 9, 7), 3), ((128, 1, 1), 0), ((256, 5, 20), 7), ((144, 17, 9), 4), ((512, 20, 5), 2),
    ((272, 8, 70), 0), ((128, 16, 2), 1),
])
def test_fused_step_bit_identical_to_f32_oracle(b200, oracle, tune, size, zc):
    tune(f"0,0,{zc},-1,0")
    w = _random_workload(b200, size, seed=sum(size) + 3)
    # a second source right at a tile corner so that the recomputed halo H' sees its m forcing
    w.sources.append(((min(127, size[0] - 1), min(7, size[1] - 1), size[2] // 2), b200.GaussianPulse(0.3, 0.4).with_amplitudes((0.5, 1.5, -2.0), (3.0, -1.0, 0.25))))
    init = random_fields(w, seed=10)
    probes = [(size[0] // 2, size[1] // 2, size[2] // 2), (0, 0, 0), (size[0] - 1, size[1] - 1, size[2] - 1), (min(128, size[0] - 1), min(8, size[1] - 1), 0)]
    steps = 5                                        # odd: the result lives in the second buffer set
    ref, s32 = oracle_run(oracle, w, steps, oracle.F32, init=init, probes=probes)
    st, series = run(b200, w, steps, exact=True, fused=True, init=init, probes=probes)
    assert st.uses_fused
    assert np.array_equal(st.read_field(b200.FieldComponent.E), ref.field(oracle.E).astype(np.float32))
    assert np.array_equal(st.read_field(b200.FieldComponent.H), ref.field(oracle.H).astype(np.float32))
    assert np.array_equal(series, s32.astype(np.float32))
    assert st.tick() == steps and st.kernel_launches < 2 * steps + 10


def test_fused_fast_mode_equals_two_kernel_path(b200, tune):
    from fdtd_b200 import configs

    tune(None)
    w = configs.config3_heterogeneous(128, 40)
    a, _ = run(b200, w, 40, exact=False, fused=True)
    b, _ = run(b200, w, 40, exact=False, fused=False)
    assert a.uses_fused and not b.uses_fused
    assert np.array_equal(a.read_field(b200.FieldComponent.E), b.read_field(b200.FieldComponent.E))
    assert np.array_equal(a.read_field(b200.FieldComponent.H), b.read_field(b200.FieldComponent.H))
    # continue stepping both one update pass at a time (reference protocol) and through reset
    for st in (a, b):
        st.step(3)
    assert np.array_equal(a.read_field(b200.FieldComponent.E), b.read_field(b200.FieldComponent.E))
    a.reset()
    assert not np.any(a.read_field(b200.FieldComponent.E)) and a.tick() == 0


def test_fused_u16_index_and_fallbacks(b200, oracle, tune):
    from fdtd_b200 import configs

    tune(None)
    rng = np.random.default_rng(12)
    n_mat = 300
    pal = [b200.Material(1.0 + 0.001 * k, 0.0, 1.0 + 0.01 * k, 0.001 * (k % 5)) for k in range(n_mat)]
    size = (128, 6, 5)
    ids = rng.integers(0, n_mat, size[0] * size[1] * size[2]).astype(np.uint32)
    w = configs.Workload("u16f", size, b200.Resolution((1, 1, 1), 0.25), b200.PhysicalConstants.REDUCED, pal, ids,
                         [((64, 3, 2), b200.GaussianPulse(1.0, 1.0).with_amplitudes((0, 0, 5.0), (0, 0, 0)))], [], 0)
    init = random_fields(w, seed=13)
    ref, _ = oracle_run(oracle, w, 4, oracle.F32, init=init)
    st, _ = run(b200, w, 4, exact=True, fused=True, init=init)
    assert st.uses_fused and st.index_bytes == 2
    assert np.array_equal(st.read_field(b200.FieldComponent.E), ref.field(oracle.E).astype(np.float32))
    assert np.array_equal(st.read_field(b200.FieldComponent.H), ref.field(oracle.H).astype(np.float32))
    assert "fused: k_step_fused<exact,u16>" in st.step_path
    # lattices the fused kernel does not support use the two-kernel path, and the handle says why
    w2 = _random_workload(b200, (36, 5, 4), seed=2)
    st2, _ = run(b200, w2, 2, exact=True, fused=True)
    assert not st2.uses_fused and st2.tick() == 2
    assert "two-kernel (vec4)" in st2.step_path and "row pitch 36 < 128" in st2.step_path


@pytest.mark.parametrize("size,cut,ranges", [((128, 9, 10), 6, "whole"), ((256, 5, 9), 4, "split"), ((128, 8, 5), 1, "whole")])
def test_fused_two_slabs_equal_one_domain(b200, tune, size, cut, ranges):
    """Two fused-step z-slab handles on one GPU, halos moved with fdtd_b200_halo_copy (the protocol distributed.py
    runs over NCCL): bit-identical to the whole-domain run, with sources and probes next to the cut."""
    from fdtd_b200 import _lib as L

    tune(None)
    w = _random_workload(b200, size, seed=41)
    nz = size[2]
    w.sources = [((size[0] // 2, size[1] // 2, cut), w.sources[0][1]), ((5, 1, cut - 1), b200.GaussianPulse(0.2, 0.5).with_amplitudes((1, 2, 3), (-1, 0.5, 2)))]
    probes = [(size[0] // 2, size[1] // 2, cut - 1), (3, 2, cut), (size[0] - 1, size[1] - 1, nz - 1)]
    init = random_fields(w, seed=14)
    steps = 5
    whole, s_whole = run(b200, w, steps, exact=True, fused=True, init=init, probes=probes)
    tab = w.source_table(steps)
    slabs = []
    for a, b in ((0, cut), (cut, nz)):
        be = b200.FdtdB200Backend(exact_arithmetic=True, fused_step=True, z_slab=(a, b))
        st = be.create_instance(w.solver_config(), w.domain()).create_state()
        assert st.uses_fused
        st.write_field(b200.FieldComponent.E, init[0][a:b], (0, 0, a), (size[0], size[1], b))
        st.write_field(b200.FieldComponent.H, init[1][a:b], (0, 0, a), (size[0], size[1], b))
        st.register_probes(probes)
        st.set_source_table(w.source_points(), tab)
        slabs.append(st)
    lo, hi = slabs
    lib = b200.lib()

    def exchange(pending):
        p = L.HALO_PENDING if pending else 0
        for d, s in ((L.HALO_RECV_EX_LOW, L.HALO_SEND_EX_TOP), (L.HALO_RECV_EY_LOW, L.HALO_SEND_EY_TOP)):
            L.check(lib.fdtd_b200_halo_copy(hi._h, d | p, lo._h, s | p))
        for d, s in ((L.HALO_RECV_EX_HIGH, L.HALO_SEND_EX_BOTTOM), (L.HALO_RECV_EY_HIGH, L.HALO_SEND_EY_BOTTOM), (L.HALO_RECV_EZ_HIGH, L.HALO_SEND_EZ_BOTTOM),
                     (L.HALO_RECV_HX_HIGH, L.HALO_SEND_HX_BOTTOM), (L.HALO_RECV_HY_HIGH, L.HALO_SEND_HY_BOTTOM)):
            L.check(lib.fdtd_b200_halo_copy(lo._h, d | p, hi._h, s | p))

    L.check(lib.fdtd_b200_halo_copy(lo._h, L.HALO_RECV_MAT_HIGH, hi._h, L.HALO_SEND_MAT_BOTTOM))
    exchange(False)
    for _ in range(steps):
        if ranges == "split" and cut >= 3:
            lo.fused_update(cut - 1, cut); hi.fused_update(0, 1)
            exchange(True)
            lo.fused_update(0, cut - 1); hi.fused_update(1, nz - cut)
        else:
            lo.fused_update(0, cut); hi.fused_update(0, nz - cut)
            exchange(True)
        for st in slabs:
            st.fused_commit(); st.end_step()
    for comp in (b200.FieldComponent.E, b200.FieldComponent.H):
        got = np.concatenate([lo.read_field(comp), hi.read_field(comp)], axis=0)
        assert np.array_equal(got, whole.read_field(comp))
    assert np.array_equal(lo.read_probes(0, steps) + hi.read_probes(0, steps), s_whole)


@pytest.mark.parametrize("size,cuts", [((128, 9, 12), (4, 9)), ((256, 5, 9), (4,)), ((128, 8, 7), (1, 3)), ((128, 6, 6), (2, 4))])
def test_p2p_slab_steps_equal_one_domain(b200, tune, size, cuts):
    """The peer-memory transport (fdtd_b200_p2p_*) with all slab handles in ONE process on one GPU (blobs carry raw pointers
    then): boundary planes first on the high-priority stream, stores into the neighbours' ghost planes, device-side flags.
    Two and three slabs (a middle slab has both neighbours; 1- and 2-plane slabs have no interior launch), sources and probes
    at the cuts, a reset + refresh in the middle: bit-identical to the whole-domain fused run."""
    tune(None)
    w = _random_workload(b200, size, seed=43)
    nz = size[2]
    c0 = cuts[0]
    w.sources = [((size[0] // 2, size[1] // 2, c0), w.sources[0][1]), ((5, 1, c0 - 1), b200.GaussianPulse(0.2, 0.5).with_amplitudes((1, 2, 3), (-1, 0.5, 2)))]
    probes = [(size[0] // 2, size[1] // 2, c0 - 1), (3, 2, c0), (size[0] - 1, size[1] - 1, nz - 1)]
    init = random_fields(w, seed=15)
    steps = 6
    tab = w.source_table(steps)
    bounds = list(zip((0,) + tuple(cuts), tuple(cuts) + (nz,)))

    def load(st, a, b):
        st.write_field(b200.FieldComponent.E, init[0][a:b], (0, 0, a), (size[0], size[1], b))
        st.write_field(b200.FieldComponent.H, init[1][a:b], (0, 0, a), (size[0], size[1], b))
        st.register_probes(probes)
        st.set_source_table(w.source_points(), tab)

    slabs = []
    for a, b in bounds:
        st = b200.FdtdB200Backend(exact_arithmetic=True, fused_step=True, z_slab=(a, b)).create_instance(w.solver_config(), w.domain()).create_state()
        assert st.uses_fused and not st.p2p_connected
        slabs.append(st)
    blobs = [st.p2p_export() for st in slabs]
    for k, st in enumerate(slabs):
        st.p2p_connect(blobs[k - 1] if k > 0 else None, blobs[k + 1] if k + 1 < len(slabs) else None)
        assert st.p2p_connected
    for rnd in range(2):   # second round: reset, reload, refresh — the flag sequence keeps counting
        whole, s_whole = run(b200, w, steps, exact=True, fused=True, init=init, probes=probes)
        for st, (a, b) in zip(slabs, bounds):
            if rnd:
                st.reset()
            load(st, a, b)
        for st in slabs:
            st.p2p_refresh()
        for _ in range(steps):          # one step per handle and turn: a single host thread must keep the slabs in lockstep
            for st in slabs:
                st.p2p_step(1)
        for st in slabs:
            st.sync()
            assert st.tick() == steps
        for comp in (b200.FieldComponent.E, b200.FieldComponent.H):
            got = np.concatenate([st.read_field(comp) for st in slabs], axis=0)
            assert np.array_equal(got, whole.read_field(comp))
        assert np.array_equal(sum(st.read_probes(0, steps) for st in slabs), s_whole)


def test_p2p_argument_errors(b200, tune):
    from fdtd_b200 import _lib as L

    tune(None)
    w = _random_workload(b200, (128, 8, 8), seed=3)
    plain = b200.FdtdB200Backend(z_slab=(0, 4), fused_step=False).create_instance(w.solver_config(), w.domain()).create_state()
    with pytest.raises(L.Fdtd_b200Error) as ei:
        plain.p2p_export()                                   # not a fused-step handle
    assert "FDTD_B200_IN_PLACE_STEP" in str(ei.value)
    a = b200.FdtdB200Backend(fused_step=True, z_slab=(0, 4)).create_instance(w.solver_config(), w.domain()).create_state()
    b = b200.FdtdB200Backend(fused_step=True, z_slab=(5, 8)).create_instance(w.solver_config(), w.domain()).create_state()
    with pytest.raises(L.Fdtd_b200Error):
        a.p2p_step(1)                                        # not connected
    ba, bb = a.p2p_export(), b.p2p_export()
    with pytest.raises(L.Fdtd_b200Error):
        a.p2p_connect(None, bb)                              # slabs are not adjacent
    with pytest.raises(L.Fdtd_b200Error):
        a.p2p_connect(None, None)                            # a neighbour blob is required above
    with pytest.raises(L.Fdtd_b200Error):
        a.p2p_connect(None, b"\0" * L.P2P_BLOB_BYTES)


@pytest.mark.parametrize("n_materials,limit", [(200, 0), (300, 0), (1500, 0), (300, 64)])
def test_fused_step_with_large_coefficient_tables(b200, oracle, tune, n_materials, limit):
    """Coefficient table staged in shared memory as long as two CTAs fit on an SM (8- and 16-bit indices), read through
    L1 beyond that (and when FDTD_B200_TUNE caps it): bit-identical to the fp32 oracle either way."""
    from fdtd_b200 import configs

    tune(f"0,0,0,-1,0,{limit}")
    size = (256, 12, 9)
    rng = np.random.default_rng(n_materials)
    pal = [b200.Material.VACUUM] + [b200.Material(1.0 + 0.01 * k, 0.001 * (k % 5), 1.0 + 0.002 * (k % 11), 0.0005 * (k % 3)) for k in range(1, n_materials)]
    ids = rng.integers(0, n_materials, size[0] * size[1] * size[2]).astype(np.uint32)
    ids[:n_materials] = np.arange(n_materials, dtype=np.uint32)
    src = b200.ContinousWave(0.3, 0.07).with_amplitudes((1.0, -2.0, 3.0), (0.5, 0.25, -0.125))
    w = configs.Workload("tables", size, b200.Resolution((1.0, 0.5, 2.0), 0.2), b200.PhysicalConstants.REDUCED, pal, ids,
                         [((128, 6, 4), src)], [], 0)
    init = random_fields(w, seed=2)
    ref, _ = oracle_run(oracle, w, 6, oracle.F32, init=init)
    st, _ = run(b200, w, 6, exact=True, fused=True, init=init)
    assert st.uses_fused
    assert np.array_equal(st.read_field(b200.FieldComponent.E), ref.field(oracle.E).astype(np.float32))
    assert np.array_equal(st.read_field(b200.FieldComponent.H), ref.field(oracle.H).astype(np.float32))


def test_a_missing_neighbour_poisons_the_handle_instead_of_yielding_stale_fields(b200, tune, monkeypatch):
    """a flag wait that gives up must be sticky and fatal.  Two slabs in one process; the lower one is
    stepped twice while the upper one never takes its first step, so the lower slab's second boundary launch waits for a push
    that does not come.  After the (shortened) timeout every entry point that steps, reads or writes state on that handle fails
    with ECUDA — no fields or probe rows computed on stale ghost planes are handed out — and the healthy handle still works."""
    from fdtd_b200 import _lib as L

    tune(None)
    monkeypatch.setenv("FDTD_B200_P2P_TIMEOUT_MS", "150")
    w = _random_workload(b200, (128, 8, 8), seed=3)
    probes = [(5, 1, 2)]
    a = b200.FdtdB200Backend(z_slab=(0, 4)).create_instance(w.solver_config(), w.domain()).create_state()
    b = b200.FdtdB200Backend(z_slab=(4, 8)).create_instance(w.solver_config(), w.domain()).create_state()
    ba, bb = a.p2p_export(), b.p2p_export()
    a.p2p_connect(None, bb)
    b.p2p_connect(ba, None)
    with pytest.raises(L.Fdtd_b200Error) as ei:
        a.p2p_step(1)                                        # nothing pushed yet: stale ghost planes are refused up front
    assert ei.value.code == L.ESTATE and "p2p_refresh" in str(ei.value)
    for st in (a, b):
        st.register_probes(probes)
        st.p2p_refresh()
    a.p2p_step(1)                                            # waits for b's refresh push: fine
    a.sync()
    with pytest.raises(L.Fdtd_b200Error) as ei:
        a.p2p_step(4)                                        # several steps per call cannot work with both slabs on one host thread
    assert ei.value.code == L.EINVAL
    a.p2p_step(1)                                            # waits for b's push of step 1, which never comes
    for call in (a.sync, lambda: a.read_field(b200.FieldComponent.E), lambda: a.read_probes(1, 1), lambda: a.p2p_step(1),
                 a.p2p_refresh, lambda: a.write_field(b200.FieldComponent.E, np.zeros((4, 8, 128, 3), np.float32))):
        with pytest.raises(L.Fdtd_b200Error) as ei:
            call()
        assert ei.value.code == L.ECUDA and "timed out waiting for a neighbour's halo flag" in str(ei.value)
    b.sync()                                                 # the other handle is not affected
    assert b.tick() == 0 and b.read_field(b200.FieldComponent.H).shape == (4, 8, 128, 3)
