"""GPU tests of the L2-resident multi-step kernel (csrc/kernels_resident.cuh): `fdtd_b200_step(h, n)` as ONE persistent
launch on lattices whose state stays in L2 — the lattices the reference application really runs (100x100x1,
cem-app/src/composer/mod.rs:841-869; 64^3, BASELINE configs[0]).  Same bars as every other kernel generation: exact mode
bit-identical to the fp32 oracle, fast mode within 1e-5 rel-L2 of the f64 oracle and bit-identical to the streaming kernels."""
import numpy as np
import pytest

from helpers import b200_run, b200_state, oracle_run, random_fields, rel_l2
from test_parity_gpu import _random_workload

pytestmark = pytest.mark.gpu


@pytest.fixture
def resident(monkeypatch):
    monkeypatch.setenv("FDTD_B200_RESIDENT", "1")
    monkeypatch.delenv("FDTD_B200_TUNE", raising=False)


def test_long_run_on_the_apps_lattice_stays_bit_identical(b200, oracle, resident):
    """2000 steps of the application's 100x100x1 configuration (cem-app/src/composer/mod.rs:841-869) in one call — 4000 cluster
    barriers between CTAs that exchange their cells through distributed shared memory — against the fp32 oracle, bit for bit."""
    from fdtd_b200 import configs

    src = b200.GaussianPulse(0.05, 0.01).with_amplitudes((0.0, 0.0, 50.0), (0.0, 0.0, 0.0))
    w = configs.Workload("app", (100, 100, 1), b200.Resolution((0.01, 0.01, 0.01), 0.001), b200.PhysicalConstants.REDUCED,
                         [b200.Material(eletrical_conductivity=10.0)], None, [((50, 50, 0), src)], [], 0)
    probes = [(60, 50, 0), (50, 60, 0), (0, 0, 0), (99, 99, 0)]
    steps = 2000
    ref, s32 = oracle_run(oracle, w, steps, oracle.F32, probes=probes)
    _, st, series = b200_run(b200, w, steps, exact=True, probes=probes, fused=None)
    assert "one cluster" in st.step_path
    assert np.array_equal(series, s32.astype(np.float32)) and np.any(series)
    assert np.array_equal(st.read_field(b200.FieldComponent.E), ref.field(oracle.E).astype(np.float32))
    assert np.array_equal(st.read_field(b200.FieldComponent.H), ref.field(oracle.H).astype(np.float32))


def _fields(st, b200):
    return st.read_field(b200.FieldComponent.E), st.read_field(b200.FieldComponent.H)


@pytest.mark.parametrize("size,kind", [
    ((100, 100, 1), "cluster"),      # the app's lattice: pitch 128 -> 12800 cells -> one cluster of 13 CTAs
    ((32, 16, 12), "cluster"), ((7, 5, 6), "cluster"), ((1, 1, 33), "cluster"), ((33, 1, 1), "cluster"), ((4, 4, 4), "cluster"),
    ((128, 16, 8), "cluster"),       # exactly 16 CTAs
    ((64, 64, 64), "grid"), ((132, 40, 30), "grid"), ((516, 33, 5), "grid"), ((48, 48, 20), "grid"),
])
def test_resident_exact_mode_is_bit_identical_to_f32_oracle(b200, oracle, resident, size, kind):
    w = _random_workload(b200, size, seed=sum(size))
    nx, ny, nz = size
    w.sources = w.sources + [((0, 0, 0), b200.GaussianPulse(0.4, 0.6).with_amplitudes((1, 2, 3), (-1, 0.5, 2))),
                             ((nx - 1, ny - 1, nz - 1), b200.GaussianPulse(0.9, 0.3).with_amplitudes((0.5, 0, -3), (2, -0.5, 1)))]
    probes = [(nx // 2, ny // 2, nz // 2), (0, 0, 0), (nx - 1, ny - 1, nz - 1), (nx // 3, ny // 2, nz // 2)]
    init = random_fields(w, seed=1)
    steps = 12
    ref, s32 = oracle_run(oracle, w, steps, oracle.F32, init=init, probes=probes)
    _, st, series = b200_run(b200, w, steps, exact=True, init=init, probes=probes, fused=None)
    kernel = "k_steps_cluster" if kind == "cluster" else "k_steps_resident"
    assert st.step_path.startswith(f"resident: {kernel}<exact,u8>") and ("one cluster" if kind == "cluster" else "cooperative grid") in st.step_path, st.step_path
    assert st.kernel_launches <= 6, "n steps are one launch"       # + the uploads' helper kernels
    e, h = _fields(st, b200)
    assert st.tick() == steps == ref.tick and st.time() == ref.time
    assert np.array_equal(e, ref.field(oracle.E).astype(np.float32))
    assert np.array_equal(h, ref.field(oracle.H).astype(np.float32))
    assert np.array_equal(series, s32.astype(np.float32))


@pytest.mark.parametrize("size", [(100, 100, 1), (64, 64, 64), (128, 24, 20)])
def test_resident_fast_mode_within_tolerance_and_equal_to_the_streaming_kernels(b200, oracle, resident, monkeypatch, size):
    w = _random_workload(b200, size, seed=5)
    probes = [(size[0] // 2 + 1, size[1] // 2, size[2] // 2)]
    init = random_fields(w, seed=2)
    steps = 40
    ref, s64 = oracle_run(oracle, w, steps, oracle.F64, init=init, probes=probes)
    _, st, series = b200_run(b200, w, steps, init=init, probes=probes, fused=None)
    assert st.step_path.startswith("resident:")
    e, h = _fields(st, b200)
    assert rel_l2(e, ref.field(oracle.E)) <= 1e-5 and rel_l2(h, ref.field(oracle.H)) <= 1e-5 and rel_l2(series, s64) <= 1e-5
    monkeypatch.setenv("FDTD_B200_RESIDENT", "0")
    _, st2, series2 = b200_run(b200, w, steps, init=init, probes=probes, fused=None)
    assert not st2.step_path.startswith("resident:")
    e2, h2 = _fields(st2, b200)
    assert np.array_equal(e, e2) and np.array_equal(h, h2) and np.array_equal(series, series2)


def test_resident_u16_index_true_division_and_batched_vs_per_step(b200, oracle, resident):
    from fdtd_b200 import configs

    rng = np.random.default_rng(3)
    n_mat = 300
    pal = [b200.Material(1.0 + 0.001 * k, 0.0, 1.0 + 0.01 * k, 0.001 * (k % 5)) for k in range(n_mat)]
    size = (40, 12, 10)
    ids = rng.integers(0, n_mat, size[0] * size[1] * size[2]).astype(np.uint32)
    src = b200.GaussianPulse(1.0, 1.0).with_amplitudes((0, 0, 5.0), (1.0, 0, 0))
    w = configs.Workload("u16r", size, b200.Resolution((0.3, 0.7, 1.1), 0.11), b200.PhysicalConstants.REDUCED, pal, ids,
                         [((20, 6, 5), src), ((20, 6, 5), src)], [], 0)          # two sources on one cell: the later wins
    probes = [(21, 6, 5), (3, 3, 3)]
    init = random_fields(w, seed=13)
    steps = 9
    ref, s32 = oracle_run(oracle, w, steps, oracle.F32, init=init, probes=probes)
    _, st, series = b200_run(b200, w, steps, exact=True, init=init, probes=probes, fused=None)
    assert st.step_path.startswith("resident: k_steps_") and "<exact-div,u16>" in st.step_path, st.step_path
    e, h = _fields(st, b200)
    assert np.array_equal(e, ref.field(oracle.E).astype(np.float32)) and np.array_equal(h, ref.field(oracle.H).astype(np.float32))
    assert np.array_equal(series, s32.astype(np.float32))
    # the reference protocol (one update pass per step, probe row read back after every step) gives the same bits
    inst, st2 = b200_state(b200, w, exact=True, fused=None)
    st2.write_field(b200.FieldComponent.E, init[0])
    st2.write_field(b200.FieldComponent.H, init[1])
    st2.register_probes(probes)
    rows = []
    t = 0.0
    for s in range(steps):
        up = inst.begin_update(st2)
        for p, f in w.sources:
            up.set_forcing(p, f.evaluate(t))
        up.finish()
        rows.append(st2.read_probes(s, 1)[0])
        t += w.resolution.temporal
    e2, h2 = _fields(st2, b200)
    assert np.array_equal(e, e2) and np.array_equal(h, h2) and np.array_equal(np.stack(rows), series)
    # mixed: resident steps, then a reset, then steps again
    st2.reset()
    assert st2.tick() == 0 and not np.any(st2.read_field(b200.FieldComponent.E))


def test_resident_is_not_used_where_it_does_not_apply(b200, resident, monkeypatch):
    w = _random_workload(b200, (128, 16, 12), seed=9)
    _, slab = b200_state(b200, w, z_slab=(0, 6), fused=None)
    assert not slab.step_path.startswith("resident:") and "z-slab" in slab.step_path
    w.pml = b200.PmlProfiles.graded(w.solver_config(), 3)
    _, st = b200_state(b200, w, fused=None)
    assert not st.step_path.startswith("resident:")
    st.step(2)
    assert st.tick() == 2
    monkeypatch.setenv("FDTD_B200_RESIDENT_MAX_CELLS", "1000")
    w2 = _random_workload(b200, (128, 16, 12), seed=9)
    _, st3 = b200_state(b200, w2, fused=None)
    assert not st3.step_path.startswith("resident:") and "on-chip budget" in st3.step_path


@pytest.mark.parametrize("n_probes,expect", [(40, "one cluster"), (700, "cooperative grid")])
def test_cluster_regime_stages_probe_rows_in_shared_memory(b200, oracle, resident, n_probes, expect):
    """The one-cluster kernel keeps the fields in distributed shared memory and stages its probe rows there until the launch ends;
    the steps per launch are bounded by the staging area (40 probes in one CTA: 17 steps per launch), and a CTA with more probe
    entries than the area holds sends the lattice to the cooperative-grid kernel.  Duplicated probe cells included."""
    size = (32, 16, 12)
    w = _random_workload(b200, size, seed=77)
    rng = np.random.default_rng(5)
    cells = rng.integers(0, 32 * 16, n_probes)              # all inside the first plane = the first CTA (512 cells)
    probes = [(int(c % 32), int((c // 32) % 16), int(c // 512)) for c in cells] + [(31, 15, 11), (31, 15, 11)]
    init = random_fields(w, seed=3)
    steps = 40
    ref, s32 = oracle_run(oracle, w, steps, oracle.F32, init=init, probes=probes)
    _, st, series = b200_run(b200, w, steps, exact=True, init=init, probes=probes, fused=None)
    assert expect in st.step_path, st.step_path
    assert np.array_equal(series, s32.astype(np.float32))
    assert np.array_equal(st.read_field(b200.FieldComponent.E), ref.field(oracle.E).astype(np.float32))
    assert np.array_equal(st.read_field(b200.FieldComponent.H), ref.field(oracle.H).astype(np.float32))
