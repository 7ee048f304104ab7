"""TEST INFRASTRUCTURE.  An independent numpy (array-slicing) formulation of the reference update
(cem-solver/src/fdtd/cpu/mod.rs:416-515, cpu/util.rs:168-194) on a z-slab with ghost planes.  Used (a) to
cross-check the C++ oracle with a second, differently-structured restatement and (b) as a CPU "slab engine"
for the gloo test of the halo-exchange orchestration."""
from __future__ import annotations

import numpy as np


class NumpySlab:
    def __init__(self, lattice, dx, dt, coeffs_zyx4, z_begin=0, z_end=None, dtype=np.float64):
        nx, ny, nz = lattice
        self.nx, self.ny, self.nz = nx, ny, nz
        self.z_begin = z_begin
        self.z_end = nz if z_end is None else z_end
        self.n_planes = self.z_end - self.z_begin
        self.lo_open = self.z_begin > 0
        self.hi_open = self.z_end < nz
        self.dx = [dtype(v) for v in dx]
        self.dt = dt
        self.dtype = dtype
        # component arrays with ghost planes at index 0 and n+1
        self.f = {c: np.zeros((self.n_planes + 2, ny, nx), dtype=dtype) for c in ("ex", "ey", "ez", "hx", "hy", "hz")}
        self.coef = np.asarray(coeffs_zyx4, dtype=dtype)[self.z_begin:self.z_end]  # [n, ny, nx, 4] = c_a c_b d_a d_b
        self.src = {}  # local (z, y, x) -> (j, m)
        self.tick = 0
        self.time = 0.0

    # -- helpers: forward / backward difference vectors with the Neumann rule ------------------------
    def _own(self, c, z_lo, z_hi):
        return self.f[c][1 + z_lo:1 + z_hi]

    def _bwd(self, c, axis, z_lo, z_hi):
        """(F[x] - F[x - e_axis]) / dx over planes [z_lo, z_hi); zero where the neighbour is missing."""
        a = self._own(c, z_lo, z_hi)
        d = np.zeros_like(a)
        if axis == 0:
            d[:, :, 1:] = a[:, :, 1:] - a[:, :, :-1]
        elif axis == 1:
            d[:, 1:, :] = a[:, 1:, :] - a[:, :-1, :]
        else:
            lower = self.f[c][z_lo:z_hi]  # plane z-1 (ghost for z == 0)
            d[...] = a - lower
            if z_lo == 0 and not self.lo_open:
                d[0] = 0
        return d / self.dx[axis]

    def _fwd(self, c, axis, z_lo, z_hi):
        a = self._own(c, z_lo, z_hi)
        d = np.zeros_like(a)
        if axis == 0:
            d[:, :, :-1] = a[:, :, 1:] - a[:, :, :-1]
        elif axis == 1:
            d[:, :-1, :] = a[:, 1:, :] - a[:, :-1, :]
        else:
            upper = self.f[c][2 + z_lo:2 + z_hi]
            d[...] = upper - a
            if z_hi == self.n_planes and not self.hi_open:
                d[-1] = 0
        return d / self.dx[axis]

    def _src_array(self, which, z_lo, z_hi):
        out = np.zeros((3, z_hi - z_lo, self.ny, self.nx), dtype=self.dtype)
        for (z, y, x), (j, m) in self.src.items():
            if z_lo <= z < z_hi:
                out[:, z - z_lo, y, x] = j if which == "j" else m
        return out

    def update_h(self, z_lo, z_hi):
        if z_lo >= z_hi:
            return
        cx = self._bwd("ez", 1, z_lo, z_hi) - self._bwd("ey", 2, z_lo, z_hi)
        cy = self._bwd("ex", 2, z_lo, z_hi) - self._bwd("ez", 0, z_lo, z_hi)
        cz = self._bwd("ey", 0, z_lo, z_hi) - self._bwd("ex", 1, z_lo, z_hi)
        m = self._src_array("m", z_lo, z_hi)
        da, db = self.coef[z_lo:z_hi, :, :, 2], self.coef[z_lo:z_hi, :, :, 3]
        for c, curl, k in (("hx", cx, 0), ("hy", cy, 1), ("hz", cz, 2)):
            own = self._own(c, z_lo, z_hi)
            own[...] = da * own + db * (-curl - m[k] + 0)

    def update_e(self, z_lo, z_hi):
        if z_lo >= z_hi:
            return
        cx = self._fwd("hz", 1, z_lo, z_hi) - self._fwd("hy", 2, z_lo, z_hi)
        cy = self._fwd("hx", 2, z_lo, z_hi) - self._fwd("hz", 0, z_lo, z_hi)
        cz = self._fwd("hy", 0, z_lo, z_hi) - self._fwd("hx", 1, z_lo, z_hi)
        j = self._src_array("j", z_lo, z_hi)
        ca, cb = self.coef[z_lo:z_hi, :, :, 0], self.coef[z_lo:z_hi, :, :, 1]
        for c, curl, k in (("ex", cx, 0), ("ey", cy, 1), ("ez", cz, 2)):
            own = self._own(c, z_lo, z_hi)
            own[...] = ca * own + cb * (curl - j[k] + 0)

    def end_step(self):
        self.tick += 1
        self.time += self.dt
        self.src = {}

    def set_forcing(self, point, j, m):
        x, y, z = point
        if self.z_begin <= z < self.z_end:
            self.src[(z - self.z_begin, y, x)] = (np.asarray(j, dtype=self.dtype), np.asarray(m, dtype=self.dtype))

    def step(self):
        self.update_h(0, self.n_planes)
        self.update_e(0, self.n_planes)
        self.end_step()

    def field(self, which):
        comps = ("ex", "ey", "ez") if which == "E" else ("hx", "hy", "hz")
        return np.stack([self.f[c][1:-1] for c in comps], axis=-1)

    # halo tensors for the SlabStepper protocol (torch views sharing memory with the numpy arrays)
    def halo(self, name):
        import torch

        n = self.n_planes
        sel = {"send_e_top": [("ex", n), ("ey", n)], "recv_e_low": [("ex", 0), ("ey", 0)],
               "send_h_bottom": [("hx", 1), ("hy", 1)], "recv_h_high": [("hx", n + 1), ("hy", n + 1)]}[name]
        return [torch.from_numpy(self.f[c][k]) for c, k in sel]
