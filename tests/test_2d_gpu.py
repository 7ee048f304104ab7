"""GPU tests of the y-marching fused step for two-dimensional lattices (csrc/kernels_2d.cuh, nz == 1 — the shape of the reference
application's own configurations, cem-app/src/composer/mod.rs:841-869).  Same bars as every kernel generation: exact mode
bit-identical to the fp32 oracle, fast mode within 1e-5 rel-L2 of the f64 oracle and bit-identical to the z-marching fused kernel.
Sources and probes sit on the seams the kernel has: between two warps' x ranges (quad 31 = x 124..127), between two row chunks,
at the lattice corners.  (The suite runs with FDTD_B200_RESIDENT=0, so these small lattices reach the streaming kernels.)"""
import numpy as np
import pytest

from helpers import b200_run, oracle_run, random_fields, rel_l2
from test_parity_gpu import _random_workload

pytestmark = pytest.mark.gpu


def _workload(b200, size, seed):
    w = _random_workload(b200, size, seed=seed)
    nx, ny, _ = size
    g = lambda a, b, j, m: b200.GaussianPulse(a, b).with_amplitudes(j, m)
    pts = [(0, 0, 0), (nx - 1, ny - 1, 0), (min(123, nx - 1), ny // 2, 0), (min(124, nx - 1), ny // 2, 0), (min(127, nx - 1), min(5, ny - 1), 0),
           (nx // 2, min(4, ny - 1), 0)]
    w.sources = w.sources + [(p, g(0.3 + 0.1 * k, 0.5, (1.0 + k, -2.0, 3.0), (0.5, 0.25 * k, -0.125))) for k, p in enumerate(pts)]
    probes = pts + [(nx // 3, ny // 3, 0), (nx // 3, ny // 3, 0)]
    return w, probes


@pytest.mark.parametrize("size,yc", [((128, 9, 1), 0), ((256, 40, 1), 5), ((516, 33, 1), 7), ((100, 100, 1), 0), ((1024, 24, 1), 8),
                                     ((4096, 6, 1), 4), ((128, 2, 1), 1), ((252, 31, 1), 16)])
def test_2d_kernel_exact_mode_is_bit_identical_to_f32_oracle(b200, oracle, monkeypatch, size, yc):
    monkeypatch.delenv("FDTD_B200_TUNE", raising=False)
    if yc:
        monkeypatch.setenv("FDTD_B200_STEP2D_YC", str(yc))
    w, probes = _workload(b200, size, seed=sum(size))
    init = random_fields(w, seed=2)
    steps = 11
    ref, s32 = oracle_run(oracle, w, steps, oracle.F32, init=init, probes=probes)
    _, st, series = b200_run(b200, w, steps, exact=True, init=init, probes=probes, fused=None)
    assert "fused 2-D: k_step_2d<exact,u8>" in st.step_path, st.step_path
    assert np.array_equal(st.read_field(b200.FieldComponent.E), ref.field(oracle.E).astype(np.float32))
    assert np.array_equal(st.read_field(b200.FieldComponent.H), ref.field(oracle.H).astype(np.float32))
    assert np.array_equal(series, s32.astype(np.float32))
    assert st.tick() == steps and st.time() == ref.time


@pytest.mark.parametrize("size", [(512, 64, 1), (300, 50, 1)])
def test_2d_kernel_fast_mode_and_true_division(b200, oracle, monkeypatch, size):
    monkeypatch.delenv("FDTD_B200_TUNE", raising=False)
    w, probes = _workload(b200, size, seed=9)
    init = random_fields(w, seed=3)
    steps = 40
    ref, s64 = oracle_run(oracle, w, steps, oracle.F64, init=init, probes=probes)
    _, st, series = b200_run(b200, w, steps, init=init, probes=probes, fused=None)
    assert "k_step_2d<fast,u8>" in st.step_path
    e, h = st.read_field(b200.FieldComponent.E), st.read_field(b200.FieldComponent.H)
    assert rel_l2(e, ref.field(oracle.E)) <= 1e-5 and rel_l2(h, ref.field(oracle.H)) <= 1e-5 and rel_l2(series, s64) <= 1e-5
    # bit-identical to the z-marching fused kernel in fast mode
    monkeypatch.setenv("FDTD_B200_STEP2D", "0")
    _, st3, series3 = b200_run(b200, w, steps, init=init, probes=probes, fused=None)
    assert "k_step_fused" in st3.step_path and "k_step_2d" not in st3.step_path
    assert np.array_equal(e, st3.read_field(b200.FieldComponent.E)) and np.array_equal(h, st3.read_field(b200.FieldComponent.H))
    assert np.array_equal(series, series3)
    monkeypatch.delenv("FDTD_B200_STEP2D")
    # dx not a power of two: exact mode divides (update.wgsl:159,171)
    w2 = _random_workload(b200, size, seed=21, dx=(0.3, 0.7, 1.1), dt=0.11)
    init2 = random_fields(w2, seed=6)
    ref2, _ = oracle_run(oracle, w2, 9, oracle.F32, init=init2)
    _, st2, _ = b200_run(b200, w2, 9, exact=True, init=init2, fused=None)
    assert "k_step_2d<exact-div,u8>" in st2.step_path
    assert np.array_equal(st2.read_field(b200.FieldComponent.E), ref2.field(oracle.E).astype(np.float32))
    assert np.array_equal(st2.read_field(b200.FieldComponent.H), ref2.field(oracle.H).astype(np.float32))


@pytest.mark.parametrize("n_mat", [300, 700])  # 700 rows: more than the kernel stages in shared memory (k2dTableMax = 512), read from global
def test_2d_kernel_u16_index_and_per_step_protocol(b200, oracle, monkeypatch, n_mat):
    from fdtd_b200 import configs

    monkeypatch.delenv("FDTD_B200_TUNE", raising=False)
    rng = np.random.default_rng(3)
    pal = [b200.Material(1.0 + 0.001 * k, 0.0, 1.0 + 0.01 * k, 0.001 * (k % 5)) for k in range(n_mat)]
    size = (256, 20, 1)
    ids = rng.integers(0, n_mat, size[0] * size[1]).astype(np.uint32)
    src = b200.GaussianPulse(1.0, 1.0).with_amplitudes((0, 0, 5.0), (1.0, 0, 0))
    w = configs.Workload("u16-2d", size, b200.Resolution((1, 1, 1), 0.25), b200.PhysicalConstants.REDUCED, pal, ids, [((124, 10, 0), src), ((124, 10, 0), src)], [], 0)
    probes = [(125, 10, 0), (3, 3, 0)]
    init = random_fields(w, seed=13)
    steps = 8
    ref, s32 = oracle_run(oracle, w, steps, oracle.F32, init=init, probes=probes)
    inst, st, series = b200_run(b200, w, steps, exact=True, init=init, probes=probes, fused=None)
    assert "k_step_2d<exact,u16>" in st.step_path
    assert np.array_equal(st.read_field(b200.FieldComponent.E), ref.field(oracle.E).astype(np.float32))
    assert np.array_equal(series, s32.astype(np.float32))
    # the reference protocol: one update pass per step
    st.reset()
    st.write_field(b200.FieldComponent.E, init[0]); st.write_field(b200.FieldComponent.H, init[1])
    t, rows = 0.0, []
    for s in range(steps):
        up = inst.begin_update(st)
        for p, f in w.sources:
            up.set_forcing(p, f.evaluate(t))
        up.finish()
        rows.append(st.read_probes(s, 1)[0])
        t += w.resolution.temporal
    assert np.array_equal(np.stack(rows), series)
    assert np.array_equal(st.read_field(b200.FieldComponent.H), ref.field(oracle.H).astype(np.float32))
    # a CPML sends the handle back to the z-marching kernel (and it says so)
    w.pml = b200.PmlProfiles.graded(w.solver_config(), 3, axes=(0, 1))
    _, stp, _ = b200_run(b200, w, 3, fused=None)
    assert "k_step_2d" not in stp.step_path and stp.tick() == 3
