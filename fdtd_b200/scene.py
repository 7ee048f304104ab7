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
from .solver import DomainDescription, Material, Resolution

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

    # DomainDescription (lib.rs:47-68) — kept so the scene also works with a per-cell backend; not used on the B200 path
    def material(self, point) -> Material:
        raise NotImplementedError("SceneDomain is rasterised on the device (fdtd_b200_set_materials_from_scene); "
                                  "the per-cell restatement lives in oracle/scene_ref.py")
