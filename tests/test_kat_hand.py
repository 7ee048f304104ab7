"""A known-answer test that does NOT pass through oracle/fdtd_ref.hpp: three steps of a unit j_z impulse on a 4x4x4 vacuum
lattice (REDUCED constants, dx = 1, dt = 1/4), computed in exact rational arithmetic straight from the reference's update
formulas (cem-solver/src/fdtd/cpu/mod.rs:416-515; curl and Neumann rule cpu/util.rs:75-194, boundary_condition.rs:29-37) and,
for the first two steps, by hand (the literals below).  Every value is a small dyadic rational, exactly representable in f32
and f64, so the oracle (both precisions) and the CUDA kernels (every generation, exact and fast arithmetic) must reproduce it
bit for bit.  The oracle stays "parity unpinned" against reference OUTPUTS (the reference has none); this pins it, and the
kernels, against an independent derivation."""
from fractions import Fraction as F

import numpy as np
import pytest

from helpers import b200_run, oracle_run

N = 4
SRC = (1, 1, 1)
DT = F(1, 4)


def rational_reference(steps):
    """E, H as dicts {(x, y, z): [Fx, Fy, Fz]} after `steps` steps.  Vacuum, eps0 = mu0 = 1, sigma = 0:
    c_a = d_a = (1 - 0)/(1 + 0) = 1, c_b = d_b = dt/(1·(1 + 0)) = dt (fdtd/util.rs:99-128); the impulse j = (0, 0, 1) acts in
    step 0 only (sources are evaluated at the pre-step time and cleared after the step, cpu/mod.rs:370-384)."""
    cells = [(x, y, z) for z in range(N) for y in range(N) for x in range(N)]
    E = {c: [F(0)] * 3 for c in cells}
    H = {c: [F(0)] * 3 for c in cells}
    unit = ((1, 0, 0), (0, 1, 0), (0, 0, 1))

    def shifted(c, axis, d):
        return tuple(v + d * u for v, u in zip(c, unit[axis]))

    def d_backward(fld, c, comp, axis):      # (F[x] − F[x − e_axis]) / dx, the whole derivative is 0 when x_axis == 0 (Neumann, alpha = 0)
        return F(0) if c[axis] == 0 else fld[c][comp] - fld[shifted(c, axis, -1)][comp]

    def d_forward(fld, c, comp, axis):       # (F[x + e_axis] − F[x]) / dx, 0 when x_axis + 1 == n (beta = 0)
        return F(0) if c[axis] + 1 == N else fld[shifted(c, axis, +1)][comp] - fld[c][comp]

    def curl(d, fld, c):                     # (dFz/dy − dFy/dz, dFx/dz − dFz/dx, dFy/dx − dFx/dy), cpu/util.rs:126-132
        return [d(fld, c, 2, 1) - d(fld, c, 1, 2), d(fld, c, 0, 2) - d(fld, c, 2, 0), d(fld, c, 1, 0) - d(fld, c, 0, 1)]

    for s in range(steps):
        j = {SRC: [F(0), F(0), F(1)]} if s == 0 else {}
        Hn = {c: [H[c][k] + DT * (-curl(d_backward, E, c)[k]) for k in range(3)] for c in cells}                # m = 0
        E = {c: [E[c][k] + DT * (curl(d_forward, Hn, c)[k] - j.get(c, [F(0)] * 3)[k]) for k in range(3)] for c in cells}
        H = Hn
    return E, H


def as_array(fld):
    out = np.zeros((N, N, N, 3), dtype=np.float64)
    for (x, y, z), v in fld.items():
        out[z, y, x] = [float(q) for q in v]
    return out


def test_hand_computed_values_of_the_first_two_steps():
    """Step 1: H stays 0, E_z(1,1,1) = −dt·j_z = −1/4.  Step 2: backward differences of that single E_z give
    H_x(1,1,1) = +1/16, H_x(1,2,1) = −1/16, H_y(1,1,1) = −1/16, H_y(2,1,1) = +1/16, and the forward differences of those give
    E_z(1,1,1) = −1/4 + 1/4·(1/8 + 1/8) = −3/16, E_z(0,1,1) = E_z(1,0,1) = −1/64, E_z(2,1,1) = E_z(1,2,1) = −1/64."""
    E, H = rational_reference(1)
    assert E[SRC] == [0, 0, F(-1, 4)] and all(v == [0, 0, 0] for c, v in E.items() if c != SRC) and all(v == [0, 0, 0] for v in H.values())
    E, H = rational_reference(2)
    assert H[(1, 1, 1)] == [F(1, 16), F(-1, 16), 0] and H[(1, 2, 1)] == [F(-1, 16), 0, 0] and H[(2, 1, 1)] == [0, F(1, 16), 0]
    assert sum(1 for v in H.values() if any(v)) == 3
    assert E[(1, 1, 1)][2] == F(-3, 16)
    for c in ((0, 1, 1), (1, 0, 1), (2, 1, 1), (1, 2, 1)):
        assert E[c][2] == F(-1, 64), c
    # E_x, E_y appear one plane below the source through the forward z-difference: curl_x = −dH_y/dz = −(−1/16 − 0), curl_y = dH_x/dz = 1/16
    assert E[(1, 1, 0)][0] == F(1, 64) and E[(1, 1, 0)][1] == F(1, 64)


def _workload(b200):
    from fdtd_b200 import configs

    class Impulse:           # j_z = 1 during step 0 (t = 0), 0 afterwards
        def evaluate(self, t):
            return b200.SourceValues((0.0, 0.0, 1.0 if t == 0.0 else 0.0), (0.0, 0.0, 0.0))

    return configs.Workload("kat-hand", (N, N, N), b200.Resolution((1.0, 1.0, 1.0), 0.25), b200.PhysicalConstants.REDUCED,
                            [b200.Material.VACUUM], None, [(SRC, Impulse())], [], 0)


@pytest.mark.parametrize("steps", [1, 2, 3])
def test_oracle_reproduces_the_rational_reference_exactly(b200, oracle, steps):
    E, H = rational_reference(steps)
    for prec in (oracle.F64, oracle.F32):
        ref, _ = oracle_run(oracle, _workload(b200), steps, prec)
        assert np.array_equal(ref.field(oracle.E), as_array(E)) and np.array_equal(ref.field(oracle.H), as_array(H))


@pytest.mark.gpu
@pytest.mark.parametrize("exact", [True, False])
@pytest.mark.parametrize("resident", ["0", "1"])
def test_cuda_kernels_reproduce_the_rational_reference_exactly(b200, exact, resident, monkeypatch):
    monkeypatch.setenv("FDTD_B200_RESIDENT", resident)
    E, H = rational_reference(3)
    for generic in (False, True):
        _, st, _ = b200_run(b200, _workload(b200), 3, exact=exact, generic=generic, fused=None)
        assert np.array_equal(st.read_field(b200.FieldComponent.E).astype(np.float64), as_array(E))
        assert np.array_equal(st.read_field(b200.FieldComponent.H).astype(np.float64), as_array(H))
