"""TEST INFRASTRUCTURE — NOT PRODUCT CODE.

ctypes front-end to the CPU restatement (oracle/fdtd_ref.hpp) of cem-solver's FDTD
time-stepping path.  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
--impl reference legs may import this package; nothing under fdtd_b200/ does.

Parity status: "parity unpinned" against reference *outputs* (the reference is Rust and
cannot be built here, and ships no FDTD golden vectors); pinned against its two
iteration-order tests and against the cited source line by line.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_DIR = os.path.dirname(os.path.abspath(__file__))
_BUILD = os.path.join(_DIR, "_build")

E, H = 0, 1
F64, F32 = 0, 1


def build(force: bool = False) -> None:
    """Compile the two oracle libraries with oracle/Makefile (gcc only, no GPU)."""
    libs = [os.path.join(_BUILD, n) for n in ("libfdtd_ref.so", "libfdtd_ref_fast.so")]
    srcs = [os.path.join(_DIR, n) for n in ("fdtd_ref.hpp", "fdtd_ref_capi.cpp", "Makefile")]
    fresh = all(os.path.exists(l) for l in libs) and min(os.path.getmtime(l) for l in libs) >= max(
        os.path.getmtime(s) for s in srcs
    )
    if fresh and not force:
        return
    subprocess.run(["make", "-C", _DIR] + (["-B"] if force else []), check=True, capture_output=True)


def _cpu_has_avx2() -> bool:
    try:
        with open("/proc/cpuinfo") as f:
            txt = f.read()
        return " avx2 " in txt and " fma " in txt and " bmi2 " in txt
    except OSError:
        return False


_libs: dict[str, C.CDLL] = {}


def _load(fast: bool) -> C.CDLL:
    name = "libfdtd_ref_fast.so" if (fast and _cpu_has_avx2()) else "libfdtd_ref.so"
    if name in _libs:
        return _libs[name]
    path = os.path.join(_BUILD, name)
    if not os.path.exists(path):
        build()
    lib = C.CDLL(path)
    u64p = C.POINTER(C.c_uint64)
    dp = C.POINTER(C.c_double)
    lib.fdtd_ref_create.restype = C.c_void_p
    lib.fdtd_ref_create.argtypes = [C.c_int, u64p, dp, C.c_double]
    lib.fdtd_ref_destroy.argtypes = [C.c_void_p]
    lib.fdtd_ref_max_threads.restype = C.c_int
    lib.fdtd_ref_set_threads.argtypes = [C.c_int]
    lib.fdtd_ref_set_materials.restype = C.c_int
    lib.fdtd_ref_set_materials.argtypes = [C.c_void_p, dp, C.c_uint32, C.POINTER(C.c_uint32), C.c_double, C.c_double]
    lib.fdtd_ref_update_coefficients.argtypes = [C.c_double, C.c_double, C.c_double, dp, dp]
    lib.fdtd_ref_pml_coefficients.argtypes = [C.c_double] * 6 + [dp, dp]
    lib.fdtd_ref_pml_coefficients_graded.argtypes = [C.c_double] * 9 + [dp, dp]
    lib.fdtd_ref_set_pml.restype = C.c_int
    lib.fdtd_ref_set_pml.argtypes = [C.c_void_p, C.c_uint64, u64p, dp, dp]
    lib.fdtd_ref_reset.argtypes = [C.c_void_p]
    lib.fdtd_ref_begin_update.argtypes = [C.c_void_p]
    lib.fdtd_ref_set_forcing.restype = C.c_int
    lib.fdtd_ref_set_forcing.argtypes = [C.c_void_p, u64p, dp, dp]
    lib.fdtd_ref_finish.argtypes = [C.c_void_p]
    lib.fdtd_ref_tick.restype = C.c_uint64
    lib.fdtd_ref_tick.argtypes = [C.c_void_p]
    lib.fdtd_ref_time.restype = C.c_double
    lib.fdtd_ref_time.argtypes = [C.c_void_p]
    lib.fdtd_ref_read_field.argtypes = [C.c_void_p, C.c_int, C.c_int, dp]
    lib.fdtd_ref_write_field.argtypes = [C.c_void_p, C.c_int, dp]
    lib.fdtd_ref_run.restype = C.c_double
    lib.fdtd_ref_run.argtypes = [C.c_void_p, C.c_uint64, C.c_uint32, u64p, dp, C.c_uint32, u64p, dp]
    lib.fdtd_ref_project.argtypes = [
        C.c_void_p, C.c_int, C.c_int, C.POINTER(C.c_float), C.POINTER(C.c_float), C.c_int, C.c_uint32, C.c_uint32,
        C.POINTER(C.c_uint8),
    ]
    lib.fdtd_ref_iter_points.restype = C.c_uint64
    lib.fdtd_ref_iter_points.argtypes = [u64p, u64p, u64p, C.c_uint64]
    lib.fdtd_ref_strider_point.argtypes = [u64p, C.c_uint64, u64p]
    lib.fdtd_ref_index_materials.restype = C.c_uint32
    lib.fdtd_ref_index_materials.argtypes = [dp, C.c_uint64, C.POINTER(C.c_uint32), dp, C.c_uint32]
    lib.fdtd_ref_gaussian_pulse.restype = C.c_double
    lib.fdtd_ref_gaussian_pulse.argtypes = [C.c_double] * 3
    lib.fdtd_ref_continous_wave.restype = C.c_double
    lib.fdtd_ref_continous_wave.argtypes = [C.c_double] * 3
    _libs[name] = lib
    return lib


def _p(a: np.ndarray, ty):
    return a.ctypes.data_as(C.POINTER(ty))


def _u64(v) -> np.ndarray:
    return np.ascontiguousarray(v, dtype=np.uint64)


def max_threads(fast: bool = False) -> int:
    return int(_load(fast).fdtd_ref_max_threads())


def set_threads(n: int, fast: bool = False) -> None:
    _load(fast).fdtd_ref_set_threads(int(n))


def update_coefficients(dt: float, eps0: float, mu0: float, material) -> np.ndarray:
    """(c_a, c_b, d_a, d_b) in f64 for material=(mu_r, sigma_m, eps_r, sigma).  util.rs:99-128."""
    m = np.ascontiguousarray(material, dtype=np.float64)
    out = np.zeros(4, dtype=np.float64)
    _load(False).fdtd_ref_update_coefficients(dt, eps0, mu0, _p(m, C.c_double), _p(out, C.c_double))
    return out


def pml_coefficients(dt: float, eps0: float, mu0: float, sigma: float, kappa: float, a: float, normal) -> tuple[np.ndarray, np.ndarray]:
    """`PmlCoefficients::new` (pml.rs:34-56): (b[3], c[3]) = scalar b, c times `normal`."""
    n = np.ascontiguousarray(normal, dtype=np.float64)
    out = np.zeros(6, dtype=np.float64)
    _load(False).fdtd_ref_pml_coefficients(dt, eps0, mu0, sigma, kappa, a, _p(n, C.c_double), _p(out, C.c_double))
    return out[:3].copy(), out[3:].copy()


def pml_coefficients_graded(dt: float, eps0: float, mu0: float, m: float, m_a: float, sigma_max: float, kappa_max: float, a_max: float,
                            depth: float, normal) -> tuple[np.ndarray, np.ndarray]:
    """`PmlCoefficients::new_graded` (pml.rs:58-86)."""
    n = np.ascontiguousarray(normal, dtype=np.float64)
    out = np.zeros(6, dtype=np.float64)
    _load(False).fdtd_ref_pml_coefficients_graded(dt, eps0, mu0, m, m_a, sigma_max, kappa_max, a_max, depth, _p(n, C.c_double),
                                                  _p(out, C.c_double))
    return out[:3].copy(), out[3:].copy()


def iter_points(start, end) -> np.ndarray:
    """Point order of util.rs:130-209 over the half-open box [start, end)."""
    s, e = _u64(start), _u64(end)
    lib = _load(False)
    n = lib.fdtd_ref_iter_points(_p(s, C.c_uint64), _p(e, C.c_uint64), None, 0)
    out = np.zeros((int(n), 3), dtype=np.uint64)
    if n:
        lib.fdtd_ref_iter_points(_p(s, C.c_uint64), _p(e, C.c_uint64), _p(out, C.c_uint64), n)
    return out


def strider_point(size, index: int) -> tuple[int, int, int]:
    n = _u64(size)
    out = np.zeros(3, dtype=np.uint64)
    _load(False).fdtd_ref_strider_point(_p(n, C.c_uint64), int(index), _p(out, C.c_uint64))
    return tuple(int(v) for v in out)


def index_materials(per_cell: np.ndarray) -> tuple[np.ndarray, np.ndarray]:
    """First-occurrence material ids + palette from per-cell (mu_r, sigma_m, eps_r, sigma) rows."""
    pc = np.ascontiguousarray(per_cell, dtype=np.float64).reshape(-1, 4)
    ids = np.zeros(pc.shape[0], dtype=np.uint32)
    lib = _load(False)
    cap = 65536
    pal = np.zeros((cap, 4), dtype=np.float64)
    k = lib.fdtd_ref_index_materials(_p(pc, C.c_double), pc.shape[0], _p(ids, C.c_uint32), _p(pal, C.c_double), cap)
    return ids, pal[: int(k)].copy()


def gaussian_pulse(time: float, t0: float, duration: float) -> float:
    return float(_load(False).fdtd_ref_gaussian_pulse(time, t0, duration))


def continous_wave(time: float, phase: float, frequency: float) -> float:
    return float(_load(False).fdtd_ref_continous_wave(time, phase, frequency))


class RefSolver:
    """The reference CPU solver (precision=F64) or the WGSL arithmetic (precision=F32) on the host."""

    def __init__(self, size, dx, dt: float, precision: int = F64, fast: bool = False):
        self._lib = _load(fast)
        self.size = tuple(int(v) for v in size)
        self.n_cells = self.size[0] * self.size[1] * self.size[2]
        self.dt = float(dt)
        n = _u64(self.size)
        d = np.ascontiguousarray(dx, dtype=np.float64)
        self._h = self._lib.fdtd_ref_create(int(precision), _p(n, C.c_uint64), _p(d, C.c_double), self.dt)
        if not self._h:
            raise MemoryError("fdtd_ref_create failed")

    def close(self) -> None:
        if getattr(self, "_h", None):
            self._lib.fdtd_ref_destroy(self._h)
            self._h = None

    def __del__(self):
        self.close()

    def set_materials(self, palette, ids=None, eps0: float = 1.0, mu0: float = 1.0) -> None:
        pal = np.ascontiguousarray(palette, dtype=np.float64).reshape(-1, 4)
        idp = None
        if ids is not None:
            ids32 = np.ascontiguousarray(np.asarray(ids).reshape(-1), dtype=np.uint32)
            assert ids32.size == self.n_cells
            idp = _p(ids32, C.c_uint32)
        rc = self._lib.fdtd_ref_set_materials(self._h, _p(pal, C.c_double), pal.shape[0], idp, eps0, mu0)
        if rc:
            raise ValueError(f"fdtd_ref_set_materials rc={rc}")

    def set_pml(self, cells, b, c, b_h=None, c_h=None) -> None:
        """Per-cell `PmlCoefficients` (cells: linear indices, b/c: [n, 3]) — the cells whose `pml(point)` is Some.
        b_h/c_h: a second set used by the H half-step (half-cell staggered grading); None = one set, the reference's model."""
        cc = _u64(cells).reshape(-1)

        def pack(bb, cv):
            out = np.ascontiguousarray(np.concatenate([np.asarray(bb, dtype=np.float64).reshape(-1, 3),
                                                       np.asarray(cv, dtype=np.float64).reshape(-1, 3)], axis=1))
            assert out.shape[0] == cc.size
            return out

        bc = pack(b, c)
        bch = pack(b_h, c_h) if b_h is not None else None
        if self._lib.fdtd_ref_set_pml(self._h, cc.size, _p(cc, C.c_uint64), _p(bc, C.c_double), _p(bch, C.c_double) if bch is not None else None):
            raise IndexError("set_pml: cell outside the lattice")

    def set_pml_profiles(self, b_axes, c_axes, b_axes_h=None, c_axes_h=None) -> None:
        """Separable PML: per-axis profiles b_axes[a][i], c_axes[a][i] (E half-step) and optionally *_h (H half-step);
        every cell with a non-zero c on some axis in either set is a PML cell."""
        nx, ny, nz = self.size
        if b_axes_h is None:
            b_axes_h, c_axes_h = b_axes, c_axes
        z, y, x = np.meshgrid(np.arange(nz), np.arange(ny), np.arange(nx), indexing="ij")
        sets = []
        for bs, cs in ((b_axes, c_axes), (b_axes_h, c_axes_h)):
            bx, by, bz = (np.asarray(v, dtype=np.float64) for v in bs)
            cx, cy, cz = (np.asarray(v, dtype=np.float64) for v in cs)
            sets.append((bx, by, bz, cx, cy, cz))
        on = np.zeros(x.shape, dtype=bool)
        for (_, _, _, cx, cy, cz) in sets:
            on |= (cx[x] != 0) | (cy[y] != 0) | (cz[z] != 0)
        cells = ((z * ny + y) * nx + x)[on]
        packed = []
        for (bx, by, bz, cx, cy, cz) in sets:
            packed.append((np.stack([bx[x[on]], by[y[on]], bz[z[on]]], axis=1), np.stack([cx[x[on]], cy[y[on]], cz[z[on]]], axis=1)))
        self.set_pml(cells, packed[0][0], packed[0][1], packed[1][0], packed[1][1])

    def reset(self) -> None:
        self._lib.fdtd_ref_reset(self._h)

    def begin_update(self) -> None:
        self._lib.fdtd_ref_begin_update(self._h)

    def set_forcing(self, point, j, m) -> None:
        p = _u64(point)
        jj = np.ascontiguousarray(j, dtype=np.float64)
        mm = np.ascontiguousarray(m, dtype=np.float64)
        if self._lib.fdtd_ref_set_forcing(self._h, _p(p, C.c_uint64), _p(jj, C.c_double), _p(mm, C.c_double)):
            raise IndexError(f"set_forcing called with invalid point: {tuple(point)}")  # reference panics

    def finish(self) -> None:
        self._lib.fdtd_ref_finish(self._h)

    @property
    def tick(self) -> int:
        return int(self._lib.fdtd_ref_tick(self._h))

    @property
    def time(self) -> float:
        return float(self._lib.fdtd_ref_time(self._h))

    def field(self, component: int, stale: bool = False) -> np.ndarray:
        """[nz, ny, nx, 3] f64 view of E or H (fresh buffer unless stale)."""
        out = np.zeros((self.n_cells, 3), dtype=np.float64)
        self._lib.fdtd_ref_read_field(self._h, component, int(stale), _p(out, C.c_double))
        return out.reshape(self.size[2], self.size[1], self.size[0], 3)

    def write_field(self, component: int, values: np.ndarray) -> None:
        v = np.ascontiguousarray(values, dtype=np.float64).reshape(-1, 3)
        assert v.shape[0] == self.n_cells
        self._lib.fdtd_ref_write_field(self._h, component, _p(v, C.c_double))

    def run(self, n_steps: int, src_points=None, jm=None, probe_cells=None):
        """Driver loop (runner.rs:522-526).  jm: [n_steps, n_src, 6].  Returns (probe series, seconds)."""
        n_src = 0 if src_points is None else len(src_points)
        sp = _u64(src_points).reshape(-1, 3) if n_src else None
        jmv = np.ascontiguousarray(jm, dtype=np.float64).reshape(n_steps, n_src, 6) if n_src else None
        n_pr = 0 if probe_cells is None else len(probe_cells)
        pc = _u64(probe_cells) if n_pr else None
        out = np.zeros((n_steps, n_pr, 6), dtype=np.float64) if n_pr else None
        secs = self._lib.fdtd_ref_run(
            self._h, n_steps, n_src, _p(sp, C.c_uint64) if n_src else None, _p(jmv, C.c_double) if n_src else None,
            n_pr, _p(pc, C.c_uint64) if n_pr else None, _p(out, C.c_double) if n_pr else None,
        )
        return out, float(secs)

    def project(self, component: int, projection, color_map, mode: int, w: int, h: int, stale: bool = True) -> np.ndarray:
        """RGBA8 [h, w, 4].  Matrices are 4x4 row-major numpy (transposed to nalgebra's column-major here)."""
        pm = np.ascontiguousarray(np.asarray(projection, dtype=np.float32).T)
        cm = np.ascontiguousarray(np.asarray(color_map, dtype=np.float32).T)
        out = np.zeros((h, w, 4), dtype=np.uint8)
        self._lib.fdtd_ref_project(self._h, component, int(stale), _p(pm, C.c_float), _p(cm, C.c_float), mode, w, h,
                                   _p(out, C.c_uint8))
        return out
