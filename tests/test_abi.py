"""CPU tests of the C-ABI shared library: it loads, exports every symbol include/fdtd_b200.h declares, and its
host-side helpers (f64 coefficient derivation, material indexing) agree bit-for-bit with the oracle.  No compute
calls are made here; on a machine without a GPU `create` must fail loudly (no CPU fallback)."""
import ctypes as C
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def header_symbols():
    txt = open(os.path.join(ROOT, "include", "fdtd_b200.h")).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    return sorted(set(re.findall(r"\b(fdtd_b200_[a-z0-9_]+)\s*\(", txt)))


def test_library_exports_every_declared_symbol(b200):
    from fdtd_b200 import _lib

    lib = b200.lib()
    names = header_symbols()
    assert len(names) >= 30
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/fdtd_b200.h but not exported"
    assert set(names) == set(_lib.SYMBOLS), "ctypes binding and header disagree"
    assert lib.fdtd_b200_abi_version() == _lib.ABI_VERSION == 2


def test_struct_layouts(b200):
    from fdtd_b200 import _lib as L

    assert C.sizeof(L.Material) == 32 and C.sizeof(L.Source) == 32 and C.sizeof(L.CoefficientsF32) == 16
    assert C.sizeof(L.Config) == 3 * 8 + 3 * 8 + 8 + 4 + 4 + 2 * 8 + 4 + 4


def test_update_coefficients_match_oracle_bitwise(b200, oracle):
    from fdtd_b200 import _lib as L

    lib = b200.lib()
    rng = np.random.default_rng(0)
    for _ in range(200):
        m = np.abs(rng.standard_normal(4)) + np.array([0.1, 0, 0.1, 0])
        dt, eps0, mu0 = abs(rng.standard_normal()) + 0.01, 8.8541878188e-12, 1.25663706127e-6
        out = L.UpdateCoefficients()
        rc = lib.fdtd_b200_update_coefficients_new(dt, C.byref(L.PhysicalConstants(eps0, mu0)), C.byref(L.Material(*m)), C.byref(out))
        assert rc == 0
        ref = oracle.update_coefficients(dt, eps0, mu0, m)
        assert [out.c_a, out.c_b, out.d_a, out.d_b] == ref.tolist()


def test_update_coefficients_nan_is_an_error(b200):
    from fdtd_b200 import _lib as L

    lib = b200.lib()
    out = L.UpdateCoefficients()
    # perm = 0 and sigma = 0 → 0/0 = NaN: the reference asserts (util.rs:112-113)
    rc = lib.fdtd_b200_update_coefficients_new(0.25, C.byref(L.PhysicalConstants(1.0, 1.0)), C.byref(L.Material(1.0, 0.0, 0.0, 0.0)), C.byref(out))
    assert rc == L.EINVAL and b"NaN" in lib.fdtd_b200_global_error()


def test_index_materials_matches_oracle(b200, oracle):
    from fdtd_b200 import _lib as L

    lib = b200.lib()
    rng = np.random.default_rng(1)
    pal = np.array([[1, 0, 1, 0], [1, 0, 3.9, 0], [2, 0.5, 1, 0], [1, 0, 1, 10], [1, 0, 1, -0.0]], dtype=np.float64)
    lab = rng.integers(0, len(pal), 5000)
    per_cell = np.ascontiguousarray(pal[lab])
    ids = np.zeros(len(lab), dtype=np.uint32)
    pal_out = (L.Material * 16)()
    n = C.c_uint32()
    rc = lib.fdtd_b200_index_materials(per_cell.ctypes.data_as(C.POINTER(L.Material)), len(lab), ids.ctypes.data_as(C.POINTER(C.c_uint32)), pal_out, 16, C.byref(n))
    assert rc == 0
    ref_ids, ref_pal = oracle.index_materials(per_cell)
    assert n.value == len(ref_pal) == 5          # -0.0 and 0.0 conductivities are different bit patterns
    assert np.array_equal(ids, ref_ids)          # integer material indexing: bit-exact
    got_pal = np.array([[p.relative_permeability, p.magnetic_conductivity, p.relative_permittivity, p.eletrical_conductivity] for p in pal_out[: n.value]])
    assert np.array_equal(got_pal.view(np.uint64), ref_pal.view(np.uint64))
    # palette_cap too small → EINVAL, count still reported
    rc = lib.fdtd_b200_index_materials(per_cell.ctypes.data_as(C.POINTER(L.Material)), len(lab), ids.ctypes.data_as(C.POINTER(C.c_uint32)), pal_out, 2, C.byref(n))
    assert rc == L.EINVAL and n.value == 5


def test_canonicalize_matches_per_cell_indexing(b200, oracle):
    from fdtd_b200 import _lib as L

    lib = b200.lib()
    rng = np.random.default_rng(2)
    pal = np.array([[1, 0, 1, 0], [1, 0, 3.9, 0], [1, 0, 1, 0], [2, 0.5, 1, 0]], dtype=np.float64)  # label 2 duplicates label 0
    lab = rng.integers(0, 4, 3000).astype(np.uint32)
    ids = np.zeros_like(lab)
    pal_out = (L.Material * 8)()
    n = C.c_uint32()
    rc = lib.fdtd_b200_canonicalize_material_index(lab.ctypes.data_as(C.POINTER(C.c_uint32)), len(lab), pal.ctypes.data_as(C.POINTER(L.Material)), 4,
                                                   ids.ctypes.data_as(C.POINTER(C.c_uint32)), pal_out, 8, C.byref(n))
    assert rc == 0 and n.value == 3
    ref_ids, _ = oracle.index_materials(pal[lab])
    assert np.array_equal(ids, ref_ids)
    bad = lab.copy(); bad[7] = 9
    rc = lib.fdtd_b200_canonicalize_material_index(bad.ctypes.data_as(C.POINTER(C.c_uint32)), len(bad), pal.ctypes.data_as(C.POINTER(L.Material)), 4,
                                                   ids.ctypes.data_as(C.POINTER(C.c_uint32)), pal_out, 8, C.byref(n))
    assert rc == L.EINVAL


def test_memory_required(b200):
    cfg = b200.FdtdSolverConfig.from_lattice((512, 512, 512), b200.Resolution((1, 1, 1), 0.25), b200.PhysicalConstants.REDUCED)
    be = b200.FdtdB200Backend(fused_step=False)
    need = be.memory_required(cfg)
    cells = 512 ** 3
    assert need == 24 * 512 * 512 * 514 + cells + 512 * 512 + 16      # fields (+2 ghost planes) + u8 index (+1 plane)
    assert need / cells < 26                      # vs 136 B/cell (cpu/mod.rs:217-221) and 80 B/cell (wgpu/mod.rs:129-133)
    assert be.memory_required(cfg, 2) == need + cells + 512 * 512
    # the default backend runs the fused step: ping-pong field sets
    assert b200.FdtdB200Backend().memory_required(cfg) == need + 24 * 512 * 512 * 514
    assert b200.FdtdB200Backend(fused_step=True).memory_required(cfg) == need + 24 * 512 * 512 * 514
    slab = b200.FdtdB200Backend(z_slab=(128, 256), fused_step=False)
    assert slab.memory_required(cfg) == 24 * 512 * 512 * 130 + 512 * 512 * 129 + 16


def test_memory_limit_picks_the_step_mode(b200):
    """cem-app/src/solver/runner.rs:304-330: memory_required is checked against memory_limit before construction.  With two
    step modes the backend takes the fused step if its two field sets fit the limit, else the in-place step, else it bails."""
    from fdtd_b200 import _lib as L

    lib = b200.lib()
    cfg = b200.FdtdSolverConfig.from_lattice((512, 512, 512), b200.Resolution((1, 1, 1), 0.25), b200.PhysicalConstants.REDUCED)
    fused = b200.FdtdB200Backend().memory_required(cfg)
    in_place = b200.FdtdB200Backend(fused_step=False).memory_required(cfg)
    assert in_place < fused
    roomy, tight, tiny = b200.FdtdB200Backend(memory_limit=fused), b200.FdtdB200Backend(memory_limit=fused - 1), b200.FdtdB200Backend(memory_limit=in_place - 1)
    assert roomy.resolve_flags(cfg) & L.IN_PLACE_STEP == 0 and roomy.memory_required(cfg) == fused
    assert tight.resolve_flags(cfg) & L.IN_PLACE_STEP and tight.memory_required(cfg) == in_place
    with pytest.raises(MemoryError):
        tiny.resolve_flags(cfg)
    with pytest.raises(MemoryError):          # before anything touches a device, like the reference's bail
        tiny.create_instance(cfg, b200.MaterialLattice([b200.Material.VACUUM])).create_state()
    with pytest.raises(MemoryError):          # an explicit mode is not overridden, only checked
        b200.FdtdB200Backend(fused_step=True, memory_limit=fused - 1).resolve_flags(cfg)
    # the C entry point itself
    cc = b200.FdtdB200Backend()._c_config(cfg)
    flags, need = C.c_uint32(99), C.c_uint64(0)
    assert lib.fdtd_b200_choose_step_mode(C.byref(cc), 1, 0, C.byref(flags), C.byref(need)) == L.OK and flags.value == 0 and need.value == fused
    assert lib.fdtd_b200_choose_step_mode(C.byref(cc), 1, fused - 1, C.byref(flags), C.byref(need)) == L.OK
    assert flags.value == L.IN_PLACE_STEP and need.value == in_place
    assert lib.fdtd_b200_choose_step_mode(C.byref(cc), 1, 1000, C.byref(flags), C.byref(need)) == L.ENOMEM and need.value == in_place
    assert lib.fdtd_b200_choose_step_mode(C.byref(cc), 3, 0, C.byref(flags), C.byref(need)) == L.EINVAL
    # the app's canned configs (cem-app/src/composer/mod.rs:841-869): 100x100x1 under memory_limit = 200 MB -> fused (pitch 128)
    app = b200.FdtdSolverConfig.from_lattice((100, 100, 1), b200.Resolution((0.01, 0.01, 0.01), 0.001), b200.PhysicalConstants.REDUCED)
    be = b200.FdtdB200Backend(memory_limit=200_000_000)
    assert be.resolve_flags(app) & L.IN_PLACE_STEP == 0 and be.memory_required(app) == 12 * 4 * 128 * 100 * 3 + 128 * 100 * 2 + 16


def test_invalid_configs_are_rejected_without_a_device(b200):
    from fdtd_b200 import _lib as L

    lib = b200.lib()
    cfg = L.Config()
    cfg.nx, cfg.ny, cfg.nz = 4, 4, 0
    cfg.dx = (C.c_double * 3)(1, 1, 1)
    cfg.dt = 0.25
    h = C.c_void_p()
    pal = (L.Material * 1)(L.Material(1, 0, 1, 0))
    pc = L.PhysicalConstants(1, 1)
    assert lib.fdtd_b200_create(C.byref(cfg), C.byref(pc), pal, 1, None, C.byref(h)) == L.EINVAL
    cfg.nz = 4
    cfg.dx = (C.c_double * 3)(1, 0, 1)
    assert lib.fdtd_b200_create(C.byref(cfg), C.byref(pc), pal, 1, None, C.byref(h)) == L.EINVAL
    cfg.dx = (C.c_double * 3)(1, 1, 1)
    cfg.z_begin, cfg.z_end = 3, 9
    assert lib.fdtd_b200_create(C.byref(cfg), C.byref(pc), pal, 1, None, C.byref(h)) == L.EINVAL
    assert h.value is None


def test_no_cpu_fallback(b200):
    """Without a CUDA device the product path must fail loudly (ECUDA), never compute on the host."""
    from fdtd_b200 import _lib as L

    if b200.lib().fdtd_b200_device_count() > 0:
        pytest.skip("a GPU is present")
    cfg = b200.FdtdSolverConfig.from_lattice((8, 8, 8), b200.Resolution((1, 1, 1), 0.25), b200.PhysicalConstants.REDUCED)
    inst = b200.FdtdB200Backend().create_instance(cfg, b200.MaterialLattice([b200.Material.VACUUM]))
    with pytest.raises(b200.Fdtd_b200Error) as ei:
        inst.create_state()
    assert ei.value.code == L.ECUDA and "no CPU fallback" in str(ei.value)


def test_product_package_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "fdtd_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                txt = open(os.path.join(dirpath, f)).read()
                assert "import oracle" not in txt and "from oracle" not in txt and "fdtd_ref" not in txt, f


def test_memory_required_counts_the_row_pitch(b200, monkeypatch):
    """Lattices with nx >= 64 and nx % 16 != 0 are stored with padded rows (include/fdtd_b200.h, fdtd_b200_row_pitch):
    150 -> 160, 96 … 127 -> 128 (the fused tile), 509 -> 512; below 64 cells, with the generic-kernel flag or with
    FDTD_B200_NO_ROW_PADDING=1 the caller's nx is kept.  Host-side rule, no GPU needed."""
    pc = b200.PhysicalConstants.REDUCED
    res = b200.Resolution((1.0, 1.0, 1.0), 0.25)

    def need(nx, **kw):
        return b200.FdtdB200Backend(**kw).memory_required(b200.FdtdSolverConfig.from_lattice((nx, 10, 8), res, pc))

    def expect(pitch, fused=None):
        plane = pitch * 10
        fused = pitch >= 128 if fused is None else fused              # the default step mode: fused wherever eligible
        return (12 if fused else 6) * 4 * plane * 10 + plane * 9 + 16

    assert need(150) == expect(160) and need(509) == expect(512) and need(100) == expect(128) and need(97) == expect(128)
    assert need(95) == expect(96) and need(80) == expect(80)
    # rows of 96 or 112 cells are padded to the fused tile only when the lattice is beyond the on-chip kernels' budget (303 104 cells)
    assert need(96) == expect(96, fused=False) and need(112) == expect(112, fused=False)        # 96 x 10 x 8 cells: stays as it is
    big = lambda nx: b200.FdtdB200Backend().memory_required(b200.FdtdSolverConfig.from_lattice((nx, 100, 80), res, pc))
    assert big(96) == big(112) == big(128) == (12 * 4 * 128 * 100 * 82 + 128 * 100 * 81 + 16)
    assert need(160) == expect(160) and need(63) == expect(63) and need(65) == expect(80)
    assert need(150) == expect(160, fused=True)                           # pitch >= 128: two buffer sets
    assert need(150, fused_step=False) == expect(160, fused=False)        # FDTD_B200_IN_PLACE_STEP
    assert need(70, fused_step=True) == expect(80, fused=False)           # pitch < 128: the fused kernel is not eligible
    assert need(150, force_generic_kernels=True) == expect(150, fused=False)
    monkeypatch.setenv("FDTD_B200_NO_ROW_PADDING", "1")
    assert need(150) == expect(150, fused=False)                          # 150 % 4 != 0: one-thread-per-cell kernels, in place
