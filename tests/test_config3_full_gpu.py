"""BASELINE configs[2] — the bench workload itself — at FULL size through the C ABI: 512³ heterogeneous lattice (64 balls
of 4 material kinds) inside the 16-cell graded lossy layer, Gaussian point source, 8 probes, 200 steps — against
tests/golden/config3_full.npz, generated from the CPU oracle by tools/make_config3_golden.py (≈ 11 min of host time on 8
cores, hence a fixture).  Exact arithmetic must reproduce the fp32 oracle bit for bit (probe series of all 200 steps and
the SHA-256 of the full 1.6 GB fields); the default arithmetic — the one bench.py times — must stay within the north-star
tolerance of the f64 oracle (probe series, a stride-5 sample of the fields inside the box the pulse has reached, their norms)."""
import hashlib
import os
import sys

import numpy as np
import pytest

from helpers import rel_l2

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLD = os.path.join(ROOT, "tests", "golden", "config3_full.npz")
TOL_VS_F64 = 1e-5


def _run(b200, w, fused, exact, steps):
    be = b200.FdtdB200Backend(exact_arithmetic=exact, fused_step=fused, probe_ring_steps=steps)
    st = be.create_instance(w.solver_config(), w.domain()).create_state()
    st.register_probes(w.probes)
    st.set_source_table(w.source_points(), w.source_table(steps))
    st.step(steps)
    return st, st.read_probes(0, steps)


@pytest.mark.parametrize("fused", [True, False])
def test_config3_full_size_bench_workload(b200, fused):
    sys.path.insert(0, ROOT)
    from bench import build_workload

    g = np.load(GOLD)
    steps, lattice = int(g["steps"]), tuple(int(v) for v in g["lattice"])
    w = build_workload(lattice, 0, lattice[2])
    assert lattice == (512, 512, 512) and np.array_equal(g["probes"], np.array(w.probes))

    # exact arithmetic == fp32 oracle, bit for bit
    st, series = _run(b200, w, fused, True, steps)
    assert st.uses_fused == fused and st.tick() == steps and st.time() == float(g["f32_time"])
    assert np.array_equal(series, g["series_f32"])
    for name, comp in (("e", b200.FieldComponent.E), ("h", b200.FieldComponent.H)):
        f = st.read_field(comp)
        assert abs(np.linalg.norm(f.astype(np.float64).ravel()) / float(g[f"f32_{name}_norm"]) - 1.0) < 1e-6   # a readable failure first
        assert hashlib.sha256(np.ascontiguousarray(f).tobytes()).hexdigest() == str(g[f"f32_{name}_sha256"]), name
        del f
    del st

    # default arithmetic (what bench.py times) within the north-star tolerance of the f64 oracle
    st, series = _run(b200, w, fused, False, steps)
    s64 = g["series_f64"]
    assert rel_l2(series, s64) <= TOL_VS_F64
    scale = max(np.linalg.norm(s64[:, k]) for k in range(len(w.probes)))
    for k in range(len(w.probes)):
        if np.linalg.norm(s64[:, k]) > 1e-3 * scale:
            assert rel_l2(series[:, k], s64[:, k]) <= 5 * TOL_VS_F64      # every probe series individually
        else:   # 200 steps of dt = 0.25 carry the pulse 50 cells: the probes 64 cells out see its numerical precursor only
            assert np.linalg.norm(series[:, k] - s64[:, k]) <= TOL_VS_F64 * scale
    for name, comp in (("e", b200.FieldComponent.E), ("h", b200.FieldComponent.H)):
        f = st.read_field(comp)
        sample = tuple(slice(int(a), int(b), int(c)) for a, b, c in g["sample_slices"])    # (z, y, x): the box the pulse has reached
        assert rel_l2(f[sample], g[f"f64_{name}_sample"]) <= TOL_VS_F64
        assert abs(np.linalg.norm(f.astype(np.float64).ravel()) / float(g[f"f64_{name}_norm"]) - 1.0) < 1e-5
        del f
