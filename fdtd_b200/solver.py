"""Host-side mirror of the reference's solver interface for the FDTD path, over the C ABI.

Names, argument meaning and error behaviour follow cem-solver:
  SolverBackend / SolverInstance / UpdatePass / UpdatePassForcing / Time / Field / FieldView
      cem-solver/src/lib.rs:28-142
  FdtdSolverConfig, Resolution                     cem-solver/src/fdtd/mod.rs:18-93
  Material, PhysicalConstants                      cem-solver/src/material.rs:21-123
  SourceValues, GaussianPulse, ContinousWave, WithAmplitudes   cem-solver/src/source.rs:10-98
  ProjectionParameters, create_projection, begin_projection_pass, add_projection, finish
      cem-solver/src/project.rs:44-133

The reference is Rust; no Rust toolchain exists in this environment, so this Python layer plays the
part of the `cem-solver-b200` shim crate shown in INTEGRATION.md.  All arithmetic on fields happens
in the CUDA library; nothing here computes field values.
"""
from __future__ import annotations

import ctypes as C
import math
from dataclasses import dataclass, field as dc_field
from typing import Callable, Iterable, Iterator, Optional, Sequence

import numpy as np

from . import _lib as L

Point = tuple  # (x, y, z) lattice coordinates, like nalgebra::Point3<usize>


class FieldComponent:
    """cem-solver/src/lib.rs:139-142"""
    E = L.FIELD_E
    H = L.FIELD_H


@dataclass(frozen=True)
class PhysicalConstants:
    """cem-solver/src/material.rs:21-55"""
    vacuum_permittivity: float
    vacuum_permeability: float

    def speed_of_light(self) -> float:
        return (self.vacuum_permittivity * self.vacuum_permeability) ** -0.5

    def frequency_to_wavelength(self, frequency: float) -> float:
        return self.speed_of_light() / frequency


PhysicalConstants.SI = PhysicalConstants(8.8541878188e-12, 1.25663706127e-6)
PhysicalConstants.REDUCED = PhysicalConstants(1.0, 1.0)


@dataclass(frozen=True)
class Material:
    """cem-solver/src/material.rs:105-123 (field order and spelling as in the reference)."""
    relative_permeability: float = 1.0
    magnetic_conductivity: float = 0.0
    relative_permittivity: float = 1.0
    eletrical_conductivity: float = 0.0

    def as_row(self) -> tuple:
        return (self.relative_permeability, self.magnetic_conductivity, self.relative_permittivity, self.eletrical_conductivity)


Material.VACUUM = Material()


@dataclass(frozen=True)
class Resolution:
    """cem-solver/src/fdtd/mod.rs:64-93"""
    spatial: tuple
    temporal: float

    @staticmethod
    def estimate_from_min_wavelength(pc: PhysicalConstants, min_wavelength: float) -> "Resolution":
        s = min_wavelength / (9.0 * math.sqrt(3.0))
        return Resolution((s, s, s), s / (pc.speed_of_light() * math.sqrt(3.0)))

    @staticmethod
    def estimate_from_max_frequency(pc: PhysicalConstants, max_frequency: float) -> "Resolution":
        t = 1.0 / (9.0 * 3.0 * max_frequency)
        s = t * pc.speed_of_light() * math.sqrt(3.0)
        return Resolution((s, s, s), t)


@dataclass(frozen=True)
class FdtdSolverConfig:
    """cem-solver/src/fdtd/mod.rs:18-38.  `size` is the physical extent; lattice dims = ceil(size/dx).max(1)."""
    resolution: Resolution
    physical_constants: PhysicalConstants
    size: tuple

    def lattice_size(self) -> tuple:
        return tuple(max(int(math.ceil(s / d)), 1) for s, d in zip(self.size, self.resolution.spatial))

    def num_cells(self) -> int:
        n = self.lattice_size()
        return n[0] * n[1] * n[2]

    @staticmethod
    def from_lattice(lattice: Sequence[int], resolution: Resolution, pc: PhysicalConstants) -> "FdtdSolverConfig":
        """Config whose lattice_size() is exactly `lattice` (avoids ceil() of inexact ratios)."""
        size = tuple((n - 0.5) * d for n, d in zip(lattice, resolution.spatial))
        cfg = FdtdSolverConfig(resolution, pc, size)
        assert cfg.lattice_size() == tuple(lattice)
        return cfg


@dataclass(frozen=True)
class SourceValues:
    """cem-solver/src/source.rs:10-13"""
    j: tuple = (0.0, 0.0, 0.0)
    m: tuple = (0.0, 0.0, 0.0)


class _ScalarSource:
    def with_amplitudes(self, j, m=(0.0, 0.0, 0.0)) -> "WithAmplitudes":
        return WithAmplitudes(SourceValues(tuple(j), tuple(m)), self)


@dataclass(frozen=True)
class GaussianPulse(_ScalarSource):
    """cem-solver/src/source.rs:39-56"""
    time: float
    duration: float

    def evaluate(self, time: float) -> float:
        q = (time - self.time) / self.duration
        return math.exp(-(q * q))


@dataclass(frozen=True)
class ContinousWave(_ScalarSource):
    """cem-solver/src/source.rs:59-76"""
    phase: float
    frequency: float

    def evaluate(self, time: float) -> float:
        return math.cos(math.tau * self.frequency * time + self.phase)


@dataclass(frozen=True)
class WithAmplitudes:
    """cem-solver/src/source.rs:79-98"""
    amplitude: SourceValues
    inner: object

    def evaluate(self, time: float) -> SourceValues:
        v = self.inner.evaluate(time)
        return SourceValues(tuple(a * v for a in self.amplitude.j), tuple(a * v for a in self.amplitude.m))


@dataclass(frozen=True)
class GradedPml:
    """cem-solver/src/fdtd/pml.rs:13-20"""
    m: float = 3.0
    m_a: float = 1.0
    sigma_max: float = 2.5
    kappa_max: float = 1.0
    a_max: float = 0.1
    normal: tuple = (1.0, 0.0, 0.0)


@dataclass(frozen=True)
class PmlCoefficients:
    """cem-solver/src/fdtd/pml.rs:27-86: the scalar (b, c) of CE 7.102 / 7.99 times the `normal` the caller passes."""
    b: tuple
    c: tuple

    @staticmethod
    def new(resolution: "Resolution", physical_constants: "PhysicalConstants", sigma: float, kappa: float, a: float, normal) -> "PmlCoefficients":
        pc = L.PhysicalConstants(physical_constants.vacuum_permittivity, physical_constants.vacuum_permeability)
        n = (C.c_double * 3)(*normal)
        b, c = (C.c_double * 3)(), (C.c_double * 3)()
        L.check(L.lib().fdtd_b200_pml_coefficients(resolution.temporal, C.byref(pc), sigma, kappa, a, n, b, c))
        return PmlCoefficients(tuple(b), tuple(c))

    @staticmethod
    def new_graded(resolution: "Resolution", physical_constants: "PhysicalConstants", m: float, m_a: float, sigma_max: float, kappa_max: float,
                   a_max: float, depth: float, normal) -> "PmlCoefficients":
        pc = L.PhysicalConstants(physical_constants.vacuum_permittivity, physical_constants.vacuum_permeability)
        g = L.GradedPml(m, m_a, sigma_max, kappa_max, a_max)
        n = (C.c_double * 3)(*normal)
        b, c = (C.c_double * 3)(), (C.c_double * 3)()
        L.check(L.lib().fdtd_b200_pml_coefficients_graded(resolution.temporal, C.byref(pc), C.byref(g), depth, n, b, c))
        return PmlCoefficients(tuple(b), tuple(c))


class PmlProfiles:
    """Separable PML in the form the C ABI takes (fdtd_b200_set_pml): per axis the (b, c) of the E half-step and of the
    H half-step over the global coordinate along that axis.  `None` for an axis = no PML along it."""

    def __init__(self, lattice: Sequence[int]):
        self.lattice = tuple(int(v) for v in lattice)
        self.axes = [None, None, None]   # (b_e, c_e, b_h, c_h) float64 arrays

    def set_axis(self, axis: int, b_e, c_e, b_h=None, c_h=None) -> "PmlProfiles":
        n = self.lattice[axis]
        arrs = [np.ascontiguousarray(v, dtype=np.float64) for v in (b_e, c_e, b_e if b_h is None else b_h, c_e if c_h is None else c_h)]
        if any(a.shape != (n,) for a in arrs):
            raise ValueError(f"PML profiles of axis {axis} must have {n} entries")
        self.axes[axis] = tuple(arrs)
        return self

    def any(self) -> bool:
        return any(a is not None and (np.any(a[1] != 0) or np.any(a[3] != 0)) for a in self.axes)

    @staticmethod
    def graded(config: "FdtdSolverConfig", thickness: int, pml: GradedPml = GradedPml(), axes: Sequence[int] = (0, 1, 2),
               staggered: bool = True) -> "PmlProfiles":
        """`thickness` PML cells on both faces of each axis, graded with `PmlCoefficients::new_graded` over the normalised
        depth.  On this lattice H[i] sits half a cell below E[i] (backward differences, cpu/mod.rs:431-443), so with
        `staggered` the H profile is evaluated at i − ½; staggered=False is the reference's one-set-per-cell model."""
        if pml.kappa_max != 1.0:
            # κ enters b and c exactly as in pml.rs:34-86, but a coordinate stretch also divides the curl's derivative by κ,
            # which neither the reference's update (cpu/mod.rs:460-461, 502-503) nor this engine's does: refuse instead of
            # silently running a different PML than the one asked for
            raise ValueError("GradedPml.kappa_max != 1 is not supported: the update has no 1/kappa stretch on the curl")
        n = config.lattice_size()
        out = PmlProfiles(n)
        for a in axes:
            t = min(int(thickness), max((n[a] - 1) // 2, 0))
            if t <= 0:
                continue
            prof = [np.zeros(n[a]) for _ in range(4)]
            hi_if = n[a] - 1 - t
            for i in range(n[a]):
                for (bi, ci, off) in ((0, 1, 0.0), (2, 3, -0.5 if staggered else 0.0)):
                    pos = i + off
                    d = max((t - pos) / t if pos < t else 0.0, (pos - hi_if) / t if pos > hi_if else 0.0)
                    d = min(d, 1.0)
                    if d > 0.0:
                        co = PmlCoefficients.new_graded(config.resolution, config.physical_constants, pml.m, pml.m_a, pml.sigma_max,
                                                        pml.kappa_max, pml.a_max, d, (1.0, 0.0, 0.0))
                        prof[bi][i], prof[ci][i] = co.b[0], co.c[0]
            out.set_axis(a, *prof)
        return out

    @staticmethod
    def from_cells(lattice: Sequence[int], cells: dict) -> "PmlProfiles":
        """Per-cell `Option<PmlCoefficients>` answers ({point: PmlCoefficients}, the reference's `pml(point)` callback,
        cpu/mod.rs:242-255) → per-axis profiles.  The coefficient of axis a must depend on the coordinate along a only;
        where a cell has none for an axis that another cell at the same coordinate has, the non-zero one is used (the
        reference's "todo: merge pmls present at this point", runner.rs:672).  The application passes −normal
        (runner.rs:667), which flips the sign of b and c on three faces: the magnitudes are used (b ≥ 0, c ≤ 0)."""
        out = PmlProfiles(lattice)
        prof = [[np.zeros(n), np.zeros(n)] for n in out.lattice]
        for point, co in cells.items():
            for a in range(3):
                b, c = abs(float(co.b[a])), -abs(float(co.c[a]))
                if c == 0.0:
                    continue
                i = int(point[a])
                if prof[a][1][i] != 0.0 and (prof[a][0][i] != b or prof[a][1][i] != c):
                    raise ValueError(f"PML is not separable: axis {a} coordinate {i} has two different coefficient sets")
                prof[a][0][i], prof[a][1][i] = b, c
        for a in range(3):
            if np.any(prof[a][1] != 0):
                out.set_axis(a, prof[a][0], prof[a][1])
        return out


class DomainDescription:
    """cem-solver/src/lib.rs:60-67: per-point material callback and per-point `pml` callback.  The reference stores the
    PML answers and never uses them (cpu/mod.rs:424-425 "todo: use pml"); this engine runs the CPML recursion with them
    (parity unpinned) — return None everywhere (the default) for the reference's executed behaviour."""

    def material(self, point: Point) -> Material:  # pragma: no cover - interface
        raise NotImplementedError

    def pml(self, point: Point):
        return None


class MaterialLattice(DomainDescription):
    """A pre-rasterised domain: palette + per-cell id lattice ([nz, ny, nx] or flat, x fastest).  This is the
    form the C ABI takes (the per-cell callback is the reference's construction bottleneck)."""

    def __init__(self, palette: Sequence[Material], ids: Optional[np.ndarray] = None, pml: Optional["PmlProfiles"] = None):
        self.palette = list(palette)
        self.ids = None if ids is None else np.ascontiguousarray(np.asarray(ids).reshape(-1), dtype=np.uint32)
        self.pml_profiles = pml

    def material(self, point: Point) -> Material:
        raise TypeError("MaterialLattice is consumed whole by create_instance")


@dataclass
class ProjectionParameters:
    """cem-solver/src/project.rs:44-68.  Matrices are 4x4, row-major numpy (M[r, c])."""
    projection: np.ndarray
    field: int
    color_map: np.ndarray
    color_map_code: Optional[str] = None  # only the app's `value.z` snippet is recognised (runner.rs:761-780)


class FdtdB200Backend:
    """`SolverBackend<FdtdSolverConfig, Point3<usize>>` (lib.rs:28-45) for the CUDA engine."""

    def __init__(self, device: int = 0, exact_arithmetic: bool = False, force_generic_kernels: bool = False,
                 z_slab: Optional[tuple] = None, probe_ring_steps: int = 0, fused_step: Optional[bool] = None,
                 memory_limit: Optional[int] = None):
        """`fused_step`: None/True = the library's default — the fused single-pass step on ping-pong buffers wherever the
        lattice is eligible (`FdtdB200SolverState.step_path` says what runs and why); False = two half-step kernels updating
        the fields in place (FDTD_B200_IN_PLACE_STEP).  `memory_limit` (bytes; `SolverConfigCommon.memory_limit`,
        cem-app/src/solver/config.rs:27-45): with fused_step=None the fused step is used if its two field sets fit the
        limit, else the in-place step; `create_instance` raises MemoryError when neither fits, which is where the
        reference bails (cem-app/src/solver/runner.rs:304-330)."""
        self.device = int(device)
        self.flags = ((L.EXACT_ARITHMETIC if exact_arithmetic else 0) | (L.FORCE_GENERIC_KERNELS if force_generic_kernels else 0)
                      | (L.IN_PLACE_STEP if fused_step is False else 0))
        self.fused_step = fused_step
        self.memory_limit = None if memory_limit is None else int(memory_limit)
        self.z_slab = z_slab
        self.probe_ring_steps = int(probe_ring_steps)

    def _c_config(self, config: FdtdSolverConfig, flags: Optional[int] = None) -> L.Config:
        n = config.lattice_size()
        c = L.Config()
        c.nx, c.ny, c.nz = n
        c.dx = (C.c_double * 3)(*config.resolution.spatial)
        c.dt = config.resolution.temporal
        c.device = self.device
        c.flags = self.flags if flags is None else flags
        c.z_begin, c.z_end = self.z_slab if self.z_slab else (0, 0)
        c.probe_ring_steps = self.probe_ring_steps
        return c

    def memory_required(self, config: FdtdSolverConfig, index_bytes: int = 1) -> Optional[int]:
        """`SolverBackend::memory_required` (lib.rs:41-44) in the step mode this backend will use for `config`."""
        cc = self._c_config(config, self.resolve_flags(config, index_bytes, enforce=False))
        return int(L.lib().fdtd_b200_memory_required(C.byref(cc), index_bytes)) or None

    def resolve_flags(self, config: FdtdSolverConfig, index_bytes: int = 1, enforce: bool = True) -> int:
        """Config flags after the memory_limit decision (fdtd_b200_choose_step_mode): fused if it fits, else in place."""
        if self.memory_limit is None or self.fused_step is not None:
            if enforce and self.memory_limit is not None:
                cc = self._c_config(config)
                need = int(L.lib().fdtd_b200_memory_required(C.byref(cc), index_bytes))
                if need > self.memory_limit:
                    raise MemoryError(f"lattice needs {need} bytes of device memory, memory_limit is {self.memory_limit}")
            return self.flags
        cc = self._c_config(config)
        flags, need = C.c_uint32(0), C.c_uint64(0)
        rc = L.lib().fdtd_b200_choose_step_mode(C.byref(cc), index_bytes, self.memory_limit, C.byref(flags), C.byref(need))
        if rc == L.ENOMEM:
            if enforce:
                raise MemoryError(f"lattice needs {need.value} bytes of device memory, memory_limit is {self.memory_limit}")
            return self.flags | L.IN_PLACE_STEP
        L.check(rc)
        return int(flags.value)

    def create_instance(self, config: FdtdSolverConfig, domain_description: DomainDescription) -> "FdtdB200SolverInstance":
        n = config.lattice_size()
        zb, ze = self.z_slab if self.z_slab else (0, n[2])
        lib = L.lib()
        if hasattr(domain_description, "c_shapes"):
            # scene.SceneDomain: rasterised on the device when the state is created — no host lattice, no per-cell callback
            inst = FdtdB200SolverInstance(self, config, domain_description.palette, None)
            inst.scene = domain_description
            inst.pml = getattr(domain_description, "pml_profiles", None)
            return inst
        if isinstance(domain_description, MaterialLattice):
            palette = domain_description.palette
            ids = domain_description.ids
            if ids is not None and ids.size == n[0] * n[1] * n[2] and (zb, ze) != (0, n[2]):
                ids = np.ascontiguousarray(ids[zb * n[0] * n[1]: ze * n[0] * n[1]])
            if ids is not None and ids.size != n[0] * n[1] * (ze - zb):
                raise ValueError("material id lattice does not match the (slab of the) lattice")
        else:
            # the reference's loop: material(point) for every linear index in order (cpu/mod.rs:244-264),
            # then compressed into first-occurrence ids + palette by the library
            cells = n[0] * n[1] * (ze - zb)
            mats = (L.Material * cells)()
            i = 0
            pml_cells = {}
            for z in range(zb, ze):
                for y in range(n[1]):
                    for x in range(n[0]):
                        co = domain_description.pml((x, y, z))
                        if co is not None:
                            pml_cells[(x, y, z)] = co
                        m = domain_description.material((x, y, z))
                        mats[i] = L.Material(*m.as_row())
                        i += 1
            ids = np.zeros(cells, dtype=np.uint32)
            pal = (L.Material * 65536)()
            npal = C.c_uint32(0)
            L.check(lib.fdtd_b200_index_materials(mats, cells, ids.ctypes.data_as(C.POINTER(C.c_uint32)), pal, 65536, C.byref(npal)))
            palette = [Material(p.relative_permeability, p.magnetic_conductivity, p.relative_permittivity, p.eletrical_conductivity)
                       for p in pal[: npal.value]]
            inst = FdtdB200SolverInstance(self, config, palette, ids)
            if pml_cells:
                inst.pml = PmlProfiles.from_cells(n, pml_cells)
            return inst
        inst = FdtdB200SolverInstance(self, config, palette, ids)
        inst.pml = getattr(domain_description, "pml_profiles", None)
        return inst


class FdtdB200SolverInstance:
    """`SolverInstance` + `Field` + projection traits.  Holds the immutable domain; `create_state` allocates
    the device-side handle (the C ABI merges Instance and State, include/fdtd_b200.h)."""

    def __init__(self, backend: FdtdB200Backend, config: FdtdSolverConfig, palette: Sequence[Material], ids: Optional[np.ndarray]):
        self.backend = backend
        self.config = config
        self.palette = list(palette)
        self.ids = ids
        self.scene = None
        self.pml: Optional[PmlProfiles] = None
        self.lattice = config.lattice_size()

    # -- SolverInstance -------------------------------------------------------------------------
    def create_state(self) -> "FdtdB200SolverState":
        lib = L.lib()
        index_bytes = 1 if len(self.palette) + 1 <= 256 else 2   # + the pad row of padded lattices (conservative)
        cc = self.backend._c_config(self.config, self.backend.resolve_flags(self.config, index_bytes))
        pc = L.PhysicalConstants(self.config.physical_constants.vacuum_permittivity, self.config.physical_constants.vacuum_permeability)
        pal = (L.Material * len(self.palette))(*[L.Material(*m.as_row()) for m in self.palette])
        idp = None if self.ids is None else self.ids.ctypes.data_as(C.POINTER(C.c_uint32))
        handle = C.c_void_p()
        L.check(lib.fdtd_b200_create(C.byref(cc), C.byref(pc), pal, len(self.palette), idp, C.byref(handle)))
        state = FdtdB200SolverState(self, handle)
        if self.scene is not None:
            sc = self.scene
            L.check(lib.fdtd_b200_set_materials_from_scene(handle, sc.c_transform(), sc.c_shapes(), len(sc.shapes), 0), handle)
        if self.pml is not None and self.pml.any():
            state.set_pml(self.pml)
        return state

    def begin_update(self, state: "FdtdB200SolverState") -> "FdtdB200UpdatePass":
        return FdtdB200UpdatePass(self, state)

    # -- Field (lib.rs:97-110) ------------------------------------------------------------------
    def field(self, state: "FdtdB200SolverState", range_: Optional[tuple], field_component: int) -> "B200FieldView":
        lo, hi = normalize_point_bounds(range_, self.lattice)
        return B200FieldView(state, lo, hi, field_component)

    # -- projection (project.rs:75-133) ---------------------------------------------------------
    def create_projection(self, state: "FdtdB200SolverState", target: "ImageTarget", parameters: ProjectionParameters) -> "B200Projection":
        return B200Projection(target, parameters)

    def begin_projection_pass(self, state: "FdtdB200SolverState") -> "B200ProjectionPass":
        return B200ProjectionPass(state)


class FdtdB200SolverState:
    """`State: Time` (lib.rs:92-95) plus the batched entry points of the C ABI."""

    def __init__(self, instance: FdtdB200SolverInstance, handle: C.c_void_p):
        self.instance = instance
        self._h = handle
        self._lib = L.lib()
        n = instance.lattice
        self.lattice = n
        zb, ze = instance.backend.z_slab if instance.backend.z_slab else (0, n[2])
        self.z_begin, self.z_end = zb, ze

    def close(self) -> None:
        if getattr(self, "_h", None):
            self._lib.fdtd_b200_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # Time
    def time(self) -> float:
        return float(self._lib.fdtd_b200_time(self._h))

    def tick(self) -> int:
        return int(self._lib.fdtd_b200_tick(self._h))

    def set_time(self, tick: int, time: float) -> None:
        L.check(self._lib.fdtd_b200_set_time(self._h, int(tick), float(time)), self._h)

    # CPML (include/fdtd_b200.h "CPML")
    def set_pml(self, profiles: Optional[PmlProfiles]) -> None:
        dp = C.POINTER(C.c_double)
        for a in range(3):
            ax = None if profiles is None else profiles.axes[a]
            if ax is None:
                L.check(self._lib.fdtd_b200_set_pml(self._h, a, None, None, None, None), self._h)
            else:
                L.check(self._lib.fdtd_b200_set_pml(self._h, a, *[v.ctypes.data_as(dp) for v in ax]), self._h)

    def has_pml(self) -> bool:
        return bool(self._lib.fdtd_b200_has_pml(self._h))

    def pml_in_kernel(self) -> bool:
        """True when the fused step kernel advances ψ itself (no separate PML passes)."""
        return bool(self._lib.fdtd_b200_pml_in_kernel(self._h))

    def read_pml_state(self) -> np.ndarray:
        out = np.zeros(int(self._lib.fdtd_b200_pml_state_floats(self._h)), dtype=np.float32)
        L.check(self._lib.fdtd_b200_read_pml_state(self._h, out.ctypes.data_as(C.POINTER(C.c_float))), self._h)
        return out

    def write_pml_state(self, values: np.ndarray) -> None:
        v = np.ascontiguousarray(values, dtype=np.float32).reshape(-1)
        if v.size != int(self._lib.fdtd_b200_pml_state_floats(self._h)):
            raise ValueError("ψ state does not match this handle's PML")
        L.check(self._lib.fdtd_b200_write_pml_state(self._h, v.ctypes.data_as(C.POINTER(C.c_float))), self._h)

    # batched API
    def cell_index(self, point: Point) -> int:
        n = self.lattice
        x, y, z = (int(v) for v in point)
        if not (0 <= x < n[0] and 0 <= y < n[1] and 0 <= z < n[2]):
            raise IndexError(f"set_forcing called with invalid point: {point!r}")  # reference panics (cpu/mod.rs:392-396)
        return x + n[0] * (y + n[1] * z)

    def set_source_table(self, points: Sequence[Point], values: np.ndarray) -> None:
        """values: [n_steps, n_src, 6] (j xyz, m xyz) for the coming steps; evaluated by the caller in f64."""
        cells = np.array([self.cell_index(p) for p in points], dtype=np.uint64)
        v = np.ascontiguousarray(values, dtype=np.float32)
        if v.ndim != 3 or v.shape[1] != len(cells) or v.shape[2] != 6:
            raise ValueError("values must be [n_steps, n_src, 6]")
        L.check(self._lib.fdtd_b200_set_source_table(self._h, len(cells), cells.ctypes.data_as(C.POINTER(C.c_uint64)), v.shape[0],
                                                     v.ctypes.data_as(C.POINTER(C.c_float))), self._h)

    def step(self, n_steps: int = 1) -> None:
        L.check(self._lib.fdtd_b200_step(self._h, int(n_steps)), self._h)

    def update_h(self, z_lo: int, z_hi: int) -> None:
        L.check(self._lib.fdtd_b200_update_h(self._h, z_lo, z_hi), self._h)

    def update_e(self, z_lo: int, z_hi: int) -> None:
        L.check(self._lib.fdtd_b200_update_e(self._h, z_lo, z_hi), self._h)

    # NVLink peer-memory halo transport (include/fdtd_b200.h "fdtd_b200_p2p_*")
    def p2p_export(self) -> bytes:
        buf = C.create_string_buffer(L.P2P_BLOB_BYTES)
        L.check(self._lib.fdtd_b200_p2p_export(self._h, buf), self._h)
        return buf.raw

    def p2p_connect(self, lower_blob: Optional[bytes], upper_blob: Optional[bytes]) -> None:
        L.check(self._lib.fdtd_b200_p2p_connect(self._h, lower_blob, upper_blob), self._h)

    def p2p_refresh(self) -> None:
        L.check(self._lib.fdtd_b200_p2p_refresh(self._h), self._h)

    def p2p_step(self, n_steps: int = 1) -> None:
        L.check(self._lib.fdtd_b200_p2p_step(self._h, int(n_steps)), self._h)

    @property
    def p2p_connected(self) -> bool:
        return bool(self._lib.fdtd_b200_p2p_connected(self._h))

    def fused_update(self, z_lo: int, z_hi: int) -> None:
        L.check(self._lib.fdtd_b200_fused_update(self._h, z_lo, z_hi), self._h)

    def fused_commit(self) -> None:
        L.check(self._lib.fdtd_b200_fused_commit(self._h), self._h)

    def end_step(self) -> None:
        L.check(self._lib.fdtd_b200_end_step(self._h), self._h)

    def sync(self) -> None:
        L.check(self._lib.fdtd_b200_sync(self._h), self._h)

    def reset(self) -> None:
        L.check(self._lib.fdtd_b200_reset(self._h), self._h)

    def set_stream(self, cuda_stream: int) -> None:
        """Run on this cudaStream_t (0 = the legacy default stream = torch's default stream)."""
        L.check(self._lib.fdtd_b200_set_stream(self._h, C.c_void_p(int(cuda_stream))), self._h)

    def reset_stream(self) -> None:
        L.check(self._lib.fdtd_b200_reset_stream(self._h), self._h)

    def register_probes(self, points: Sequence[Point]) -> None:
        cells = np.array([self.cell_index(p) for p in points], dtype=np.uint64)
        self._n_probes = len(cells)
        L.check(self._lib.fdtd_b200_register_probes(self._h, len(cells), cells.ctypes.data_as(C.POINTER(C.c_uint64))), self._h)

    def read_probes(self, first_tick: int, n_steps: int) -> np.ndarray:
        out = np.zeros((n_steps, self._n_probes, 6), dtype=np.float32)
        L.check(self._lib.fdtd_b200_read_probes(self._h, first_tick, n_steps, out.ctypes.data_as(C.POINTER(C.c_float))), self._h)
        return out

    def halo_plane(self, which: int) -> tuple:
        p = C.c_void_p()
        n = C.c_uint64()
        L.check(self._lib.fdtd_b200_halo_plane(self._h, which, C.byref(p), C.byref(n)), self._h)
        return int(p.value), int(n.value)

    def device_field(self, c: int) -> int:
        p = C.c_void_p()
        L.check(self._lib.fdtd_b200_device_field(self._h, c, C.byref(p)), self._h)
        return int(p.value)

    @property
    def row_pitch(self) -> int:
        """Cells per stored row (>= lattice[0]; include/fdtd_b200.h, fdtd_b200_row_pitch)."""
        return int(self._lib.fdtd_b200_row_pitch(self._h))

    def read_field(self, component: int, lo: Optional[Point] = None, hi: Optional[Point] = None) -> np.ndarray:
        """float32 [bz, by, bx, 3] of the box [lo, hi) (default: this handle's whole slab)."""
        n = self.lattice
        lo = (0, 0, self.z_begin) if lo is None else tuple(int(v) for v in lo)
        hi = (n[0], n[1], self.z_end) if hi is None else tuple(int(v) for v in hi)
        lo_c = tuple(min(l, s) for l, s in zip(lo, n))
        hi_c = tuple(min(max(h, l), s) for h, l, s in zip(hi, lo, n))
        shape = (max(hi_c[2] - lo_c[2], 0), max(hi_c[1] - lo_c[1], 0), max(hi_c[0] - lo_c[0], 0), 3)
        out = np.zeros(shape, dtype=np.float32)
        a = (C.c_uint64 * 3)(*lo)
        b = (C.c_uint64 * 3)(*hi)
        L.check(self._lib.fdtd_b200_read_field(self._h, component, a, b, out.ctypes.data_as(C.POINTER(C.c_float))), self._h)
        return out

    def write_field(self, component: int, values: np.ndarray, lo: Optional[Point] = None, hi: Optional[Point] = None) -> None:
        n = self.lattice
        lo = (0, 0, self.z_begin) if lo is None else tuple(int(v) for v in lo)
        hi = (n[0], n[1], self.z_end) if hi is None else tuple(int(v) for v in hi)
        v = np.ascontiguousarray(values, dtype=np.float32)
        if v.size != 3 * (hi[0] - lo[0]) * (hi[1] - lo[1]) * (hi[2] - lo[2]):
            raise ValueError("values do not match the box")
        a = (C.c_uint64 * 3)(*lo)
        b = (C.c_uint64 * 3)(*hi)
        L.check(self._lib.fdtd_b200_write_field(self._h, component, a, b, v.ctypes.data_as(C.POINTER(C.c_float))), self._h)

    def set_profiling(self, enabled: bool) -> None:
        L.check(self._lib.fdtd_b200_set_profiling(self._h, int(enabled)), self._h)

    def profile(self, which: int) -> tuple:
        """(total device ms, launches, cells updated) of the H (0) or E (1) half-step kernels."""
        ms, n, cells = C.c_double(), C.c_uint64(), C.c_uint64()
        L.check(self._lib.fdtd_b200_get_profile(self._h, which, C.byref(ms), C.byref(n), C.byref(cells)), self._h)
        return ms.value, n.value, cells.value

    def timer_start(self) -> None:
        L.check(self._lib.fdtd_b200_timer_start(self._h), self._h)

    def timer_stop(self) -> float:
        ms = C.c_double()
        L.check(self._lib.fdtd_b200_timer_stop(self._h, C.byref(ms)), self._h)
        return ms.value

    def material_ids(self) -> np.ndarray:
        n = self.lattice
        out = np.zeros(n[0] * n[1] * (self.z_end - self.z_begin), dtype=np.uint32)
        L.check(self._lib.fdtd_b200_read_material_ids(self._h, out.ctypes.data_as(C.POINTER(C.c_uint32))), self._h)
        return out

    def coefficient_table(self) -> np.ndarray:
        n = C.c_uint32()
        L.check(self._lib.fdtd_b200_read_coefficient_table(self._h, None, 0, C.byref(n)), self._h)
        tab = (L.CoefficientsF32 * n.value)()
        L.check(self._lib.fdtd_b200_read_coefficient_table(self._h, tab, n.value, C.byref(n)), self._h)
        return np.ctypeslib.as_array(tab).view(np.float32).reshape(-1, 4).copy()

    @property
    def index_bytes(self) -> int:
        return int(self._lib.fdtd_b200_index_bytes(self._h))

    @property
    def uses_vec4(self) -> bool:
        return bool(self._lib.fdtd_b200_uses_vec4(self._h))

    @property
    def uses_fused(self) -> bool:
        return bool(self._lib.fdtd_b200_uses_fused(self._h))

    @property
    def uses_bulk(self) -> bool:
        return bool(self._lib.fdtd_b200_uses_bulk(self._h))

    @property
    def kernel_launches(self) -> int:
        return int(self._lib.fdtd_b200_kernel_launches(self._h))

    @property
    def step_path(self) -> str:
        """Which kernels `step` launches on this handle and why (fdtd_b200_step_path) — there are no silent fallbacks."""
        buf = C.create_string_buffer(512)
        self._lib.fdtd_b200_step_path(self._h, buf, len(buf))
        return buf.value.decode()

    @property
    def memory_in_use(self) -> int:
        return int(self._lib.fdtd_b200_memory_in_use(self._h))


class FdtdB200UpdatePass:
    """`UpdatePass` + `UpdatePassForcing` (lib.rs:81-90; cpu/mod.rs:363-409)."""

    def __init__(self, instance: FdtdB200SolverInstance, state: FdtdB200SolverState):
        self.state = state
        self._sources: list = []  # begin_update starts with no forcing (cpu/mod.rs:375-381)

    def set_forcing(self, point: Point, value: SourceValues) -> None:
        cell = self.state.cell_index(point)  # raises where the reference panics
        self._sources.append((cell, value))

    def finish(self) -> None:
        st = self.state
        n = len(self._sources)
        arr = (L.Source * max(n, 1))()
        for k, (cell, v) in enumerate(self._sources):
            arr[k].cell = cell
            arr[k].j = (C.c_float * 3)(*[float(x) for x in v.j])
            arr[k].m = (C.c_float * 3)(*[float(x) for x in v.m])
        L.check(st._lib.fdtd_b200_set_sources(st._h, n, arr), st._h)
        L.check(st._lib.fdtd_b200_step(st._h, 1), st._h)


def normalize_point_bounds(range_: Optional[tuple], size: Sequence[int]) -> tuple:
    """cem-solver/src/fdtd/util.rs:140-162 for `(lo, hi)` half-open boxes; None bounds are unbounded.
    `("incl", lo, hi)` expresses the reference's `lo..=hi`."""
    if range_ is None:
        lo, hi = None, None
    elif len(range_) == 3 and range_[0] == "incl":
        lo, hi = range_[1], tuple(v + 1 for v in range_[2])
    else:
        lo, hi = range_
    lo = (0, 0, 0) if lo is None else tuple(int(v) for v in lo)
    hi = tuple(size) if hi is None else tuple(int(v) for v in hi)
    hi = tuple(max(a, b) for a, b in zip(lo, hi))
    return lo, hi


class B200FieldView:
    """`FieldView` (lib.rs:129-136): values widened to f64 like wgpu/mod.rs:522."""

    def __init__(self, state: FdtdB200SolverState, lo: Point, hi: Point, component: int):
        self.state, self.lo, self.hi, self.component = state, lo, hi, component
        self._cache: Optional[np.ndarray] = None

    def _data(self) -> np.ndarray:
        if self._cache is None:
            self._cache = self.state.read_field(self.component, self.lo, self.hi).astype(np.float64)
        return self._cache

    def at(self, point: Point) -> Optional[np.ndarray]:
        n = self.state.lattice
        if not all(l <= p < h for p, l, h in zip(point, self.lo, self.hi)):
            return None
        if not all(0 <= p < s for p, s in zip(point, n)):
            return None
        v = self.state.read_field(self.component, tuple(point), tuple(p + 1 for p in point))
        return v.reshape(3).astype(np.float64)

    def array(self) -> np.ndarray:
        return self._data()

    def iter(self) -> Iterator[tuple]:
        d = self._data()
        bz, by, bx, _ = d.shape
        for z in range(bz):
            for y in range(by):
                for x in range(bx):
                    yield (self.lo[0] + x, self.lo[1] + y, self.lo[2] + z), d[z, y, x]


class ImageTarget:
    """Headless stand-in for the render-texture target (cem-render texture channel, RGBA8, size = lattice.xy;
    cem-app/src/solver/runner.rs:791-797): a host RGBA8 image."""

    def __init__(self, width: int, height: int):
        self.image = np.zeros((height, width, 4), dtype=np.uint8)

    def size(self) -> tuple:
        return self.image.shape[1], self.image.shape[0]

    def with_image_buffer(self, f) -> None:
        """`FdtdImageTarget::with_image_buffer` (project.rs:136-151): hand the renderer this target's buffer."""
        f(self.image)


_APP_Z_SNIPPET_MARK = "var x = value.z"


class B200Projection:
    def __init__(self, target: ImageTarget, parameters: ProjectionParameters, mode: Optional[int] = None):
        self.target = target
        self.parameters = parameters
        if mode is not None:
            self.mode = mode
        elif parameters.color_map_code is not None:
            if _APP_Z_SNIPPET_MARK not in parameters.color_map_code:
                raise ValueError("only the app's value.z colour-map snippet is supported (runner.rs:761-780)")
            self.mode = L.COLORMAP_APP_Z
        else:
            self.mode = L.COLORMAP_LINEAR


class B200ProjectionPass:
    """`ProjectionPass` + `ProjectionPassAdd` (project.rs:112-133).  Renders the fresh field (the reference
    renders the previous step's buffer, cpu/project.rs:107 — see DESIGN.md)."""

    def __init__(self, state: FdtdB200SolverState):
        self.state = state
        self.errors: list = []

    def add_projection(self, projection: B200Projection) -> None:
        st = self.state
        p = projection.parameters
        pm = np.ascontiguousarray(np.asarray(p.projection, dtype=np.float32).T)  # → column-major
        cm = np.ascontiguousarray(np.asarray(p.color_map, dtype=np.float32).T)
        def render(img: np.ndarray) -> None:
            L.check(st._lib.fdtd_b200_project(st._h, p.field, pm.ctypes.data_as(C.POINTER(C.c_float)), cm.ctypes.data_as(C.POINTER(C.c_float)),
                                              projection.mode, img.shape[1], img.shape[0], img.ctypes.data_as(C.c_void_p), 0), st._h)

        try:
            projection.target.with_image_buffer(render)   # ImageTarget, io.GifEncoder, io.PngSequence …
        except L.Fdtd_b200Error as e:  # collected like cpu/project.rs:185-189
            self.errors.append(e)

    def finish(self) -> None:
        if self.errors:
            raise RuntimeError(f"fdtd-b200 projection pass error: {self.errors}")
