"""`fdtd_b200_read_field_static` — the field readback that allocates nothing (its gather kernel's scratch is a file-scope
`static __device__` array of FDTD_B200_STATIC_SCRATCH_BYTES instead of a cudaMalloc'd staging buffer; no counterpart in the
reference) — against `fdtd_b200_read_field` through the C ABI: bit-identical (a
gather moves values, the tolerance of the field-readback tests is exact equality), on two calls in a row, the second smaller
than the first so that stale contents of the shared array would show.  Collected last on purpose;
passes alone in a fresh process and after the rest of the suite."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _box(lib, fn, h, comp, lo, hi):
    n = max(hi[0] - lo[0], 0) * max(hi[1] - lo[1], 0) * max(hi[2] - lo[2], 0)
    out = np.full((n, 3), np.nan, dtype=np.float32)
    a, b = (C.c_uint64 * 3)(*lo), (C.c_uint64 * 3)(*hi)
    rc = fn(h, comp, a, b, out.ctypes.data_as(C.POINTER(C.c_float)))
    return rc, out


def test_static_scratch_readback_equals_the_allocating_one_twice_in_a_row(b200):
    from fdtd_b200 import _lib as L
    from fdtd_b200 import configs

    lib = b200.lib()
    size = (100, 36, 20)                                  # rows padded to 128 cells inside the library: the box walk is not contiguous
    rng = np.random.default_rng(11)
    pal = [b200.Material.VACUUM, b200.Material(relative_permittivity=3.9), b200.Material(eletrical_conductivity=10.0)]
    ids = rng.integers(0, 3, size[0] * size[1] * size[2]).astype(np.uint32)
    src = b200.GaussianPulse(1.0, 0.7).with_amplitudes((1.0, -2.0, 3.0), (0.5, 0.25, -0.125))
    w = configs.Workload("zz", size, b200.Resolution((1.0, 0.5, 2.0), 0.2), b200.PhysicalConstants.REDUCED, pal, ids, [((50, 18, 10), src)], [], 0)
    st = b200.FdtdB200Backend().create_instance(w.solver_config(), w.domain()).create_state()
    # no write_field here: it would allocate the handle's staging buffer, and the point below is that the static call needs none
    st.set_source_table(w.source_points(), w.source_table(40))
    st.step(30)
    h = st._h
    in_use = st.memory_in_use

    # call 1: the whole lattice (72 000 cells), E and H; call 2 straight after: a small interior box — both before the allocating
    # entry point has run on this handle, so no staging buffer exists yet
    full = ((0, 0, 0), size)
    small = ((7, 3, 2), (61, 30, 17))
    got = {}
    for name, (lo, hi) in (("full", full), ("small", small)):
        for comp in (L.FIELD_E, L.FIELD_H):
            rc, out = _box(lib, lib.fdtd_b200_read_field_static, h, comp, lo, hi)
            assert rc == L.OK, lib.fdtd_b200_last_error(h)
            got[name, comp] = out
    assert st.memory_in_use == in_use, "the static entry point must not have grown the handle's staging buffer"
    for name, (lo, hi) in (("full", full), ("small", small)):
        for comp in (L.FIELD_E, L.FIELD_H):
            rc, ref = _box(lib, lib.fdtd_b200_read_field, h, comp, lo, hi)
            assert rc == L.OK
            assert not np.isnan(ref).any() and np.any(ref != 0)
            assert np.array_equal(got[name, comp], ref), (name, comp)
    assert st.memory_in_use > in_use                       # ... whereas the allocating one did

    # two more calls in a row after further steps, in the other order (small box first, then a larger one over the same array)
    st.step(4)
    for lo, hi in (small, ((0, 0, 5), (100, 36, 15))):
        rc, a = _box(lib, lib.fdtd_b200_read_field_static, h, L.FIELD_E, lo, hi)
        rc2, b = _box(lib, lib.fdtd_b200_read_field, h, L.FIELD_E, lo, hi)
        assert rc == rc2 == L.OK and np.array_equal(a, b)
    # clipping and empty boxes behave like the allocating entry point (util.rs:140-162)
    rc, a = _box(lib, lib.fdtd_b200_read_field_static, h, L.FIELD_H, (90, 30, 18), (100, 36, 20))
    rc2, b = _box(lib, lib.fdtd_b200_read_field, h, L.FIELD_H, (90, 30, 18), (100, 36, 20))
    assert rc == rc2 == L.OK and np.array_equal(a, b)
    rc, _ = _box(lib, lib.fdtd_b200_read_field_static, h, L.FIELD_E, (5, 5, 5), (5, 9, 9))
    assert rc == L.OK
    st.close()


def test_boxes_beyond_the_static_array_are_refused_with_a_clear_error(b200):
    from fdtd_b200 import _lib as L

    lib = b200.lib()
    max_cells = L.STATIC_SCRATCH_BYTES // 12
    assert max_cells == 699_050
    cfg = b200.FdtdSolverConfig.from_lattice((128, 128, 48), b200.Resolution((1, 1, 1), 0.25), b200.PhysicalConstants.REDUCED)   # 786 432 cells
    st = b200.FdtdB200Backend().create_instance(cfg, b200.MaterialLattice([b200.Material.VACUUM])).create_state()
    out = np.zeros((128 * 128 * 48, 3), dtype=np.float32)
    a, b = (C.c_uint64 * 3)(0, 0, 0), (C.c_uint64 * 3)(128, 128, 48)
    rc = lib.fdtd_b200_read_field_static(st._h, L.FIELD_E, a, b, out.ctypes.data_as(C.POINTER(C.c_float)))
    msg = lib.fdtd_b200_last_error(st._h).decode()
    assert rc == L.EINVAL and "786432 cells exceeds the static scratch array" in msg and "699050" in msg, msg
    b = (C.c_uint64 * 3)(128, 128, 42)                      # 688 128 cells: fits
    assert lib.fdtd_b200_read_field_static(st._h, L.FIELD_E, a, b, out.ctypes.data_as(C.POINTER(C.c_float))) == L.OK
    assert lib.fdtd_b200_read_field_static(st._h, 7, a, b, out.ctypes.data_as(C.POINTER(C.c_float))) == L.EINVAL
    st.close()
