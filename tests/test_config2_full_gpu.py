"""BASELINE configs[1] at FULL size through the C ABI: 256³ lattice with dielectric / lossy / magnetic inclusions and a
conductor slab, CW source, 8 probes sampled every step, 5000 steps — against tests/golden/config2_full.npz, generated from
the CPU oracle by tools/make_config2_golden.py (30 min of host time, hence a fixture).  Exact arithmetic must reproduce
the fp32 oracle bit for bit (probe series of all 5000 steps, SHA-256 of the full fields at steps 500 and 5000); the
default arithmetic must stay within the north-star tolerance of the f64 oracle."""
import hashlib
import os

import numpy as np
import pytest

from helpers import rel_l2

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "config2_full.npz")
TOL_VS_F64 = 1e-5


def _run(b200, w, fused, exact, mid, steps):
    be = b200.FdtdB200Backend(exact_arithmetic=exact, fused_step=fused, probe_ring_steps=steps)
    st = be.create_instance(w.solver_config(), w.domain()).create_state()
    st.register_probes(w.probes)
    st.set_source_table(w.source_points(), w.source_table(steps))
    st.step(mid)
    snap_mid = (st.read_field(b200.FieldComponent.E), st.read_field(b200.FieldComponent.H))
    st.step(steps - mid)
    snap_end = (st.read_field(b200.FieldComponent.E), st.read_field(b200.FieldComponent.H))
    return st, st.read_probes(0, steps), snap_mid, snap_end


@pytest.mark.parametrize("fused", [True, False])
def test_config2_full_size_5000_steps(b200, fused):
    from fdtd_b200 import configs

    g = np.load(GOLD)
    steps, mid = int(g["steps"]), int(g["mid"])
    w = configs.config2_inclusions(256, steps)
    assert tuple(g["lattice"]) == w.lattice and np.array_equal(g["probes"], np.array(w.probes))

    # exact arithmetic == fp32 oracle, bit for bit
    st, series, snap_mid, snap_end = _run(b200, w, fused, True, mid, steps)
    assert st.uses_fused == fused and st.tick() == steps and st.time() == float(g["f32_time"])
    assert np.array_equal(series, g["series_f32"])
    for tag, (e, h) in (("mid", snap_mid), ("end", snap_end)):
        for name, f in (("e", e), ("h", h)):
            assert abs(np.linalg.norm(f.astype(np.float64).ravel()) / float(g[f"f32_{name}_{tag}_norm"]) - 1.0) < 1e-6   # a readable failure first
            assert hashlib.sha256(np.ascontiguousarray(f).tobytes()).hexdigest() == str(g[f"f32_{name}_{tag}_sha256"]), (tag, name)

    # default arithmetic within the north-star tolerance of the f64 oracle (probe series and a 1-in-512 sample of the fields)
    st, series, snap_mid, snap_end = _run(b200, w, fused, False, mid, steps)
    s64 = g["series_f64"]
    assert rel_l2(series, s64) <= TOL_VS_F64
    scale = max(np.linalg.norm(s64[:, k]) for k in range(len(w.probes)))
    for k in range(len(w.probes)):
        if np.linalg.norm(s64[:, k]) > 1e-6 * scale:
            assert rel_l2(series[:, k], s64[:, k]) <= 5 * TOL_VS_F64      # every probe series individually
        else:   # the probe behind the conductor slab sees ≈ 1e-14 of the others: rounding noise, bounded absolutely
            assert np.linalg.norm(series[:, k] - s64[:, k]) <= TOL_VS_F64 * scale
    for tag, (e, h) in (("mid", snap_mid), ("end", snap_end)):
        assert rel_l2(e[::8, ::8, ::8], g[f"f64_e_{tag}_sample"]) <= TOL_VS_F64
        assert rel_l2(h[::8, ::8, ::8], g[f"f64_h_{tag}_sample"]) <= TOL_VS_F64
        assert abs(np.linalg.norm(e.astype(np.float64).ravel()) / float(g[f"f64_e_{tag}_norm"]) - 1.0) < 1e-5
