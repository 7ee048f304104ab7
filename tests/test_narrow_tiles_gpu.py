"""GPU tests of the 64-cell tile of the fused step kernel (k_step_fused<…, BX = 16, BY = 16>: two thread rows per warp, 15 core rows),
which plain fused launches take when the last 128-cell tile column of a row would be at most half full (fdtd_b200.cu:
use_narrow_tiles).  Same bars as every kernel generation: exact mode bit-identical to the fp32 oracle, fast mode bit-identical to
the 128-cell tile and within 1e-5 rel-L2 of the f64 oracle.  Sources and probes sit on the tile seams (x = 63 | 64, the last core
row 14 | the recomputed row 15 of a tile, chunk seams) and at the lattice corners."""
import numpy as np
import pytest

from helpers import b200_run, oracle_run, random_fields, rel_l2
from test_parity_gpu import _random_workload

pytestmark = pytest.mark.gpu


def _workload(b200, size, seed, **kw):
    w = _random_workload(b200, size, seed=seed, **kw)
    nx, ny, nz = size
    g = lambda a, b, j, m: b200.GaussianPulse(a, b).with_amplitudes(j, m)
    pts = [(0, 0, 0), (nx - 1, ny - 1, nz - 1), (63, min(14, ny - 1), nz // 2), (64, min(15, ny - 1), nz // 2), (min(127, nx - 1), min(29, ny - 1), 0),
           (min(128, nx - 1), min(30, ny - 1), nz - 1), (nx // 2, ny // 2, nz // 2)]
    w.sources = w.sources + [(p, g(0.3 + 0.1 * k, 0.5, (1.0 + k, -2.0, 3.0), (0.5, 0.25 * k, -0.125))) for k, p in enumerate(pts)]
    return w, pts + [(nx // 3, ny // 3, nz // 3)]


@pytest.mark.parametrize("size,zc", [((192, 33, 9), 0), ((320, 20, 12), 4), ((160, 47, 6), 0), ((136, 16, 5), 2), ((576, 18, 7), 0)])
def test_narrow_tiles_exact_mode_is_bit_identical_to_f32_oracle(b200, oracle, monkeypatch, size, zc):
    monkeypatch.setenv("FDTD_B200_TUNE", f"0,0,{zc}") if zc else monkeypatch.delenv("FDTD_B200_TUNE", raising=False)
    monkeypatch.delenv("FDTD_B200_NARROW_TILES", raising=False)
    w, probes = _workload(b200, size, seed=sum(size))
    init = random_fields(w, seed=2)
    steps = 9
    ref, s32 = oracle_run(oracle, w, steps, oracle.F32, init=init, probes=probes)
    _, st, series = b200_run(b200, w, steps, exact=True, init=init, probes=probes, fused=None)
    assert "k_step_fused<exact,u8>" in st.step_path and "64-cell tiles" in st.step_path, st.step_path
    assert np.array_equal(st.read_field(b200.FieldComponent.E), ref.field(oracle.E).astype(np.float32))
    assert np.array_equal(st.read_field(b200.FieldComponent.H), ref.field(oracle.H).astype(np.float32))
    assert np.array_equal(series, s32.astype(np.float32))


def test_narrow_tiles_fast_mode_equals_the_wide_tile_and_the_choice_is_reported(b200, oracle, monkeypatch):
    monkeypatch.delenv("FDTD_B200_TUNE", raising=False)
    size = (300, 40, 20)      # row pitch 304 = 2.4 tiles of 128: padded rows (xpad) AND narrow tiles
    w, probes = _workload(b200, size, seed=5)
    init = random_fields(w, seed=3)
    steps = 30
    ref, s64 = oracle_run(oracle, w, steps, oracle.F64, init=init, probes=probes)
    _, st, series = b200_run(b200, w, steps, init=init, probes=probes, fused=None)
    assert "64-cell tiles" in st.step_path and "xpad" in st.step_path, st.step_path
    e, h = st.read_field(b200.FieldComponent.E), st.read_field(b200.FieldComponent.H)
    assert rel_l2(e, ref.field(oracle.E)) <= 1e-5 and rel_l2(h, ref.field(oracle.H)) <= 1e-5 and rel_l2(series, s64) <= 1e-5
    monkeypatch.setenv("FDTD_B200_NARROW_TILES", "0")
    _, st2, series2 = b200_run(b200, w, steps, init=init, probes=probes, fused=None)
    assert "k_step_fused" in st2.step_path and "64-cell tiles" not in st2.step_path
    assert np.array_equal(e, st2.read_field(b200.FieldComponent.E)) and np.array_equal(h, st2.read_field(b200.FieldComponent.H))
    assert np.array_equal(series, series2)
    # a lattice whose rows fill their 128-cell tiles keeps them; FDTD_B200_NARROW_TILES=1 forces the narrow ones (u16 index, true division)
    monkeypatch.delenv("FDTD_B200_NARROW_TILES")
    w3 = _random_workload(b200, (256, 24, 10), seed=4)
    _, st3, _ = b200_run(b200, w3, 2, fused=None)
    assert "k_step_fused" in st3.step_path and "64-cell tiles" not in st3.step_path
    monkeypatch.setenv("FDTD_B200_NARROW_TILES", "1")
    from fdtd_b200 import configs
    rng = np.random.default_rng(3)
    pal = [b200.Material(1.0 + 0.001 * k, 0.0, 1.0 + 0.01 * k, 0.001 * (k % 5)) for k in range(300)]
    ids = rng.integers(0, 300, 256 * 24 * 10).astype(np.uint32)
    src = b200.GaussianPulse(1.0, 1.0).with_amplitudes((0, 0, 5.0), (1.0, 0, 0))
    w4 = configs.Workload("u16-narrow", (256, 24, 10), b200.Resolution((0.3, 0.7, 1.1), 0.11), b200.PhysicalConstants.REDUCED, pal, ids,
                          [((64, 15, 5), src), ((63, 14, 4), src)], [], 0)
    init4 = random_fields(w4, seed=6)
    ref4, _ = oracle_run(oracle, w4, 7, oracle.F32, init=init4)
    _, st4, _ = b200_run(b200, w4, 7, exact=True, init=init4, fused=None)
    assert "k_step_fused<exact-div,u16>" in st4.step_path and "64-cell tiles" in st4.step_path, st4.step_path
    assert np.array_equal(st4.read_field(b200.FieldComponent.E), ref4.field(oracle.E).astype(np.float32))
    assert np.array_equal(st4.read_field(b200.FieldComponent.H), ref4.field(oracle.H).astype(np.float32))
