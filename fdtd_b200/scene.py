"""Scene → material lattice: the application side of `DomainDescription::material`.

In the reference every lattice point is pushed through `CoordinateTransformations::transform_point_from_solver_to_world`
(f64 4×4, then cast to f32; cem-app/src/solver/runner.rs:873-918) and looked up in the scene's BVH; the first collider that
contains the point supplies the material, else the solver's default material (runner.rs:616-631).  That is one host
callback per cell on one thread (cem-solver/src/fdtd/cpu/mod.rs:244-264).  Here the scene is a short ordered list of
primitives (parry3d Ball / Cuboid / Cylinder with an isometry) handed to the CUDA rasteriser
(`fdtd_b200_rasterize_scene`, `fdtd_b200_set_materials_from_scene`), which writes the index lattice on the device.

"First" = list order here; the reference's BVH traversal order is unspecified (parity unpinned), as is parry's
exact floating-point containment predicate — oracle/scene_ref.py states the predicate this library uses.
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass, field
from typing import Optional, Sequence

import numpy as np

from . import _lib as L
from .solver import DomainDescription, GradedPml, Material, PhysicalConstants, PmlCoefficients, PmlProfiles, Resolution

BALL, CUBOID, CYLINDER = 0, 1, 2


def quaternion_to_matrix(q: Sequence[float]) -> np.ndarray:
    """Unit quaternion (w, x, y, z) → 3×3 rotation, f64."""
    w, x, y, z = (float(v) for v in q)
    n = (w * w + x * x + y * y + z * z) ** 0.5
    w, x, y, z = w / n, x / n, y / n, z / n
    return np.array([[1 - 2 * (y * y + z * z), 2 * (x * y - w * z), 2 * (x * z + w * y)],
                     [2 * (x * y + w * z), 1 - 2 * (x * x + z * z), 2 * (y * z - w * x)],
                     [2 * (x * z - w * y), 2 * (y * z + w * x), 1 - 2 * (x * x + y * y)]], dtype=np.float64)


@dataclass(frozen=True)
class Shape:
    """A collider with its `GlobalTransform` isometry (translation + rotation) and the material of its entity."""
    kind: int
    params: tuple                    # ball (r,) | cuboid (hx, hy, hz) | cylinder (half_height, radius) — axis = local Y
    material: Material
    translation: tuple = (0.0, 0.0, 0.0)
    rotation: tuple = (1.0, 0.0, 0.0, 0.0)   # unit quaternion w, x, y, z

    @staticmethod
    def ball(radius: float, material: Material, translation=(0.0, 0.0, 0.0)) -> "Shape":
        return Shape(BALL, (float(radius),), material, tuple(translation))

    @staticmethod
    def cuboid(half_extents, material: Material, translation=(0.0, 0.0, 0.0), rotation=(1.0, 0.0, 0.0, 0.0)) -> "Shape":
        return Shape(CUBOID, tuple(float(v) for v in half_extents), material, tuple(translation), tuple(rotation))

    @staticmethod
    def cylinder(half_height: float, radius: float, material: Material, translation=(0.0, 0.0, 0.0), rotation=(1.0, 0.0, 0.0, 0.0)) -> "Shape":
        return Shape(CYLINDER, (float(half_height), float(radius)), material, tuple(translation), tuple(rotation))

    def world_to_local_rotation(self) -> np.ndarray:
        """Row-major f32 3×3: the inverse (transpose) of the entity's rotation."""
        return np.ascontiguousarray(quaternion_to_matrix(self.rotation).T, dtype=np.float32)


@dataclass(frozen=True)
class PmlRegion:
    """An entity carrying a `GradedPml` component and a cuboid collider (the only shape the application gives one,
    cem-app/src/composer/presets.rs:119-143).  `pml(point)` of the reference (solver/runner.rs:633-675) finds the entities whose
    AABB contains the point, casts a ray from the point along `pml.normal` against the collider (non-solid, at most the AABB's
    diagonal far) and feeds the time of impact — the distance to the face the normal points at — as `depth` into
    `PmlCoefficients::new_graded` together with −normal.  parry3d's f32 ray/cuboid routine is third-party (parry3d 0.25,
    un-vendored): restated here as the slab method in f32 — parity unpinned."""
    half_extents: tuple
    pml: GradedPml
    translation: tuple = (0.0, 0.0, 0.0)
    rotation: tuple = (1.0, 0.0, 0.0, 0.0)

    def aabb(self) -> tuple:
        r = np.abs(quaternion_to_matrix(self.rotation)).astype(np.float32)
        e = r @ np.asarray(self.half_extents, dtype=np.float32)
        c = np.asarray(self.translation, dtype=np.float32)
        return c - e, c + e

    def cast_ray(self, origin: np.ndarray, direction: np.ndarray, max_toi: float) -> Optional[float]:
        """Time of impact of origin + t·direction with the cuboid's surface (solid = false: a ray that starts inside hits
        the face it leaves through), None beyond max_toi or on a miss."""
        rot = quaternion_to_matrix(self.rotation).T.astype(np.float32)
        o = rot @ (np.asarray(origin, dtype=np.float32) - np.asarray(self.translation, dtype=np.float32))
        d = rot @ np.asarray(direction, dtype=np.float32)
        h = np.asarray(self.half_extents, dtype=np.float32)
        tmin, tmax = np.float32(0.0), np.float32(max_toi)
        inside = bool(np.all(np.abs(o) <= h))
        t_exit = np.float32(np.inf)
        for a in range(3):
            if d[a] == 0.0:
                if abs(o[a]) > h[a]:
                    return None
                continue
            inv = np.float32(1.0) / d[a]
            t0, t1 = (-h[a] - o[a]) * inv, (h[a] - o[a]) * inv
            if t0 > t1:
                t0, t1 = t1, t0
            tmin = max(tmin, t0)
            t_exit = min(t_exit, t1)
            if tmin > min(t_exit, tmax):
                return None
        toi = t_exit if inside else tmin
        return float(toi) if np.isfinite(toi) and 0.0 <= toi <= max_toi else None


@dataclass
class CoordinateTransformations:
    """runner.rs:873-931."""
    transform_from_solver_to_world: np.ndarray      # 4×4 f64
    transform_from_world_to_solver: np.ndarray
    lattice_size: tuple

    @staticmethod
    def for_fdtd(resolution: Resolution, lattice_size: Sequence[int], rotation: Sequence[float] = (1.0, 0.0, 0.0, 0.0),
                 aabb_mins: Sequence[float] = (0.0, 0.0, 0.0)) -> "CoordinateTransformations":
        scale = np.diag([*resolution.spatial, 1.0]).astype(np.float64)
        inv_scale = np.diag([*(1.0 / v for v in resolution.spatial), 1.0]).astype(np.float64)
        iso = np.eye(4, dtype=np.float64)
        iso[:3, :3] = quaternion_to_matrix(rotation)
        iso[:3, 3] = np.asarray(aabb_mins, dtype=np.float32).astype(np.float64)   # `aabb.mins.coords.cast::<f64>()` of an f32 Aabb
        inv_iso = np.eye(4, dtype=np.float64)
        inv_iso[:3, :3] = iso[:3, :3].T
        inv_iso[:3, 3] = -iso[:3, :3].T @ iso[:3, 3]
        return CoordinateTransformations(iso @ scale, inv_scale @ inv_iso, tuple(int(v) for v in lattice_size))

    def transform_point_from_solver_to_world(self, point: Sequence[int]) -> np.ndarray:
        m = self.transform_from_solver_to_world
        x, y, z = (float(v) for v in point)
        return np.array([((m[r, 0] * x + m[r, 1] * y) + m[r, 2] * z) + m[r, 3] for r in range(3)], dtype=np.float64).astype(np.float32)

    def transform_point_from_world_to_solver(self, point: Sequence[float]) -> Optional[tuple]:
        p = self.transform_from_world_to_solver @ np.array([*np.asarray(point, dtype=np.float32).astype(np.float64), 1.0])
        r = np.sign(p[:3]) * np.floor(np.abs(p[:3]) + 0.5)   # f64::round: half away from zero
        if np.any(r < 0) or np.any(r >= np.asarray(self.lattice_size)):
            return None
        return tuple(int(v) for v in r)


@dataclass
class SceneDomain(DomainDescription):
    """A `DomainDescription` the backend rasterises on the device instead of calling `material(point)` per cell."""
    shapes: list
    transformations: CoordinateTransformations
    default_material: Material = Material.VACUUM
    palette: list = field(default_factory=list)     # filled by __post_init__: [default] + distinct shape materials (list order)
    labels: list = field(default_factory=list)
    # entities with a GradedPml component (runner.rs:594-604); need the solver's resolution and constants for the coefficients
    pml_regions: list = field(default_factory=list)
    resolution: Optional[Resolution] = None
    physical_constants: Optional[PhysicalConstants] = None
    _pml_override: Optional[PmlProfiles] = field(default=None, repr=False)   # profiles assigned by the caller win over the regions

    def __post_init__(self):
        self.palette, self.labels = [self.default_material], []
        for s in self.shapes:
            if s.material not in self.palette:
                self.palette.append(s.material)
            self.labels.append(self.palette.index(s.material))

    def c_shapes(self):
        arr = (L.Shape * max(len(self.shapes), 1))()
        for i, (s, lab) in enumerate(zip(self.shapes, self.labels)):
            arr[i].kind, arr[i].label = s.kind, lab
            p = list(s.params) + [0.0] * (3 - len(s.params))
            arr[i].params = (C.c_float * 3)(*p)
            arr[i].rot = (C.c_float * 9)(*s.world_to_local_rotation().reshape(-1))
            arr[i].translation = (C.c_float * 3)(*s.translation)
        return arr

    def c_transform(self):
        return (C.c_double * 16)(*np.ascontiguousarray(self.transformations.transform_from_solver_to_world, dtype=np.float64).reshape(-1))

    def rasterize(self, device: int = 0, z_begin: int = 0, z_end: Optional[int] = None) -> np.ndarray:
        """Labels (indices into `palette`) of planes [z_begin, z_end), u32, x fastest — computed on the GPU."""
        n = self.transformations.lattice_size
        z_end = n[2] if z_end is None else z_end
        out = np.empty((z_end - z_begin) * n[1] * n[0], dtype=np.uint32)
        dims = (C.c_uint64 * 3)(*n)
        L.check(L.lib().fdtd_b200_rasterize_scene(device, dims, z_begin, z_end, self.c_transform(), self.c_shapes(), len(self.shapes), 0,
                                                  out.ctypes.data_as(C.POINTER(C.c_uint32))))
        return out

    def pml(self, point) -> Optional[PmlCoefficients]:
        """`DomainDescription::pml` as the application answers it (runner.rs:633-675): the first PML entity whose AABB holds the
        point and whose collider the ray along its normal hits supplies graded coefficients (no merging: runner.rs:672)."""
        if not self.pml_regions:
            return None
        if self.resolution is None or self.physical_constants is None:
            raise ValueError("SceneDomain with PML regions needs resolution and physical_constants")
        p = self.transformations.transform_point_from_solver_to_world(point)
        for reg in self.pml_regions:
            lo, hi = reg.aabb()
            if np.any(p < lo) or np.any(p > hi):
                continue
            normal = np.asarray(reg.pml.normal, dtype=np.float32)
            toi = reg.cast_ray(p, normal, float(np.linalg.norm((hi - lo).astype(np.float32))))
            if toi is None:
                continue
            g = reg.pml
            return PmlCoefficients.new_graded(self.resolution, self.physical_constants, g.m, g.m_a, g.sigma_max, g.kappa_max, g.a_max,
                                              toi, tuple(-float(v) for v in normal.astype(np.float64)))
        return None

    @property
    def pml_profiles(self) -> Optional[PmlProfiles]:
        """The separable per-axis form the device runs (fdtd_b200_set_pml), without a per-cell sweep: every region must be a
        slab of the lattice — normal along a lattice axis, covering the lattice's whole cross-section — so that its answer
        depends on the coordinate along the normal only; the profile is `pml(point)` sampled along the lattice's centre lines.
        Anything else raises (as `PmlProfiles.from_cells` does for a non-separable cell set)."""
        if self._pml_override is not None or not self.pml_regions:
            return self._pml_override
        n = self.transformations.lattice_size
        centre = [v // 2 for v in n]
        corners = [(x, y, z) for x in (0, n[0] - 1) for y in (0, n[1] - 1) for z in (0, n[2] - 1)]
        cells = {}
        for a in range(3):
            for i in range(n[a]):
                pt = list(centre)
                pt[a] = i
                co = self.pml(pt)
                if co is None:
                    continue
                cells[tuple(pt)] = co
                # separability: the same coordinate along a must answer the same at the far corners of the cross-section
                for cpt in corners:
                    q = list(cpt)
                    q[a] = i
                    cq = self.pml(q)
                    if cq is None or not np.allclose(cq.c, co.c, rtol=1e-5, atol=0.0) or not np.allclose(cq.b, co.b, rtol=1e-5, atol=0.0):
                        raise ValueError(f"PML region is not a slab of the lattice: axis {a} coordinate {i} answers differently at {tuple(q)} "
                                         f"and {tuple(pt)} (the device runs separable per-axis profiles)")
        return PmlProfiles.from_cells(n, cells) if cells else None

    @pml_profiles.setter
    def pml_profiles(self, profiles: Optional[PmlProfiles]) -> None:
        self._pml_override = profiles

    # DomainDescription (lib.rs:47-68) — kept so the scene also works with a per-cell backend; not used on the B200 path
    def material(self, point) -> Material:
        raise NotImplementedError("SceneDomain is rasterised on the device (fdtd_b200_set_materials_from_scene); "
                                  "the per-cell restatement lives in oracle/scene_ref.py")
