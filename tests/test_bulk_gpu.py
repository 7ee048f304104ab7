"""GPU tests of the TMA bulk-copy staged kernels (csrc/kernels_bulk.cuh, FDTD_B200_TUNE variant 1): they must be
bit-identical to the fp32 oracle in exact mode, on clipped tiles, x/y halos, short z-chunks and z-slabs."""
import os

import numpy as np
import pytest

from helpers import b200_run, b200_state, oracle_run, random_fields
from test_parity_gpu import _random_workload

pytestmark = pytest.mark.gpu


@pytest.fixture
def tune():
    old = os.environ.get("FDTD_B200_TUNE")

    def set_(v):
        os.environ["FDTD_B200_TUNE"] = v

    yield set_
    if old is None:
        os.environ.pop("FDTD_B200_TUNE", None)
    else:
        os.environ["FDTD_B200_TUNE"] = old


@pytest.mark.parametrize("size,tune_str", [
    ((512, 6, 9), "128,2,4,1,3"),      # full-row tile, y clipped to 6 rows = 3 tiles, zc=4 → 3 chunks (last short)
    ((512, 5, 7), "128,2,3,1,4"),      # odd ny: last tile has one row
    ((512, 7, 6), "128,1,8,1,4"),
    ((512, 9, 5), "128,4,2,1,2"),
    ((512, 9, 11), "128,4,16,1,3"),
    ((528, 4, 6), "128,2,3,1,3"),      # two x tiles: the second has 16 columns and a left/right halo
    ((1040, 3, 5), "128,2,5,1,3"),     # three x tiles
    ((48, 10, 9), "32,4,4,1,4"),       # lattice narrower than the tile
    ((160, 6, 8), "32,4,3,1,4"),       # 32-wide CTAs, two x tiles (128 + 32 columns)
    ((16, 1, 1), "32,4,4,1,4"),
    ((64, 1, 40), "32,4,7,1,4"),
])
def test_bulk_kernels_bit_identical_to_f32_oracle(b200, oracle, tune, size, tune_str):
    tune(tune_str)
    w = _random_workload(b200, size, seed=sum(size) + 1)
    init = random_fields(w, seed=8)
    steps = 7
    ref, _ = oracle_run(oracle, w, steps, oracle.F32, init=init)
    _, st, _ = b200_run(b200, w, steps, exact=True, init=init)
    assert np.array_equal(st.read_field(b200.FieldComponent.E), ref.field(oracle.E).astype(np.float32))
    assert np.array_equal(st.read_field(b200.FieldComponent.H), ref.field(oracle.H).astype(np.float32))


def test_bulk_kernels_probes_sources_and_fast_mode(b200, oracle, tune):
    from fdtd_b200 import configs

    w = configs.config2_inclusions(64, 60)
    ref32, s32 = oracle_run(oracle, w, 60, oracle.F32, probes=w.probes)
    tune("0,0,0,0,0")
    _, st_reg, s_reg = b200_run(b200, w, 60, exact=False, probes=w.probes)
    tune("32,4,16,1,4")
    _, st_x, s_x = b200_run(b200, w, 60, exact=True, probes=w.probes)
    _, st_f, s_f = b200_run(b200, w, 60, exact=False, probes=w.probes)
    assert np.array_equal(s_x, s32.astype(np.float32))
    assert np.array_equal(st_x.read_field(b200.FieldComponent.E), ref32.field(oracle.E).astype(np.float32))
    # fast arithmetic is the same instruction sequence in both kernel families
    assert np.array_equal(st_f.read_field(b200.FieldComponent.E), st_reg.read_field(b200.FieldComponent.E))
    assert np.array_equal(st_f.read_field(b200.FieldComponent.H), st_reg.read_field(b200.FieldComponent.H))
    assert np.array_equal(s_f, s_reg)


def test_bulk_kernels_on_slabs(b200, tune):
    from fdtd_b200 import _lib as L

    tune("128,2,3,1,3")
    size, cut, steps = (512, 6, 10), 4, 6
    w = _random_workload(b200, size, seed=31)
    init = random_fields(w, seed=9)
    _, whole, _ = b200_run(b200, w, steps, init=init)
    _, lo = b200_state(b200, w, z_slab=(0, cut))
    _, hi = b200_state(b200, w, z_slab=(cut, size[2]))
    lib = b200.lib()
    tab = w.source_table(steps)
    for st, (a, b) in ((lo, (0, cut)), (hi, (cut, size[2]))):
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
        lo.update_h(0, 1); lo.update_h(1, cut); hi.update_h(0, 1); hi.update_h(1, size[2] - cut)
        h_down()
        lo.update_e(cut - 1, cut); lo.update_e(0, cut - 1); hi.update_e(0, size[2] - cut)
        e_up()
        lo.end_step(); hi.end_step()
    for comp in (b200.FieldComponent.E, b200.FieldComponent.H):
        got = np.concatenate([lo.read_field(comp), hi.read_field(comp)], axis=0)
        assert np.array_equal(got, whole.read_field(comp))
