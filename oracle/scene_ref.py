"""TEST INFRASTRUCTURE — CPU restatement of the scene → material-lattice step, the checker for the CUDA rasteriser
(fdtd_b200/csrc/kernels_scene.cuh).  Only tests/ and __graft_entry__.smoke() may import this.

What it restates: the reference fills its lattice by calling `DomainDescription::material(point)` for every linear index
in order (cem-solver/src/fdtd/cpu/mod.rs:244-264); the application's implementation maps the lattice point to world
coordinates with an f64 4×4 matrix, casts to f32 (cem-app/src/solver/runner.rs:909-918), and takes the material of the first
collider containing the point, else the default (runner.rs:616-631).  Containment is parry3d 0.25 `PointQuery::contains_point`
(cem-scene/src/spatial/traits.rs:79-86) on Ball / Cuboid / Cylinder; parry is an un-vendored dependency, so the exact
floating-point predicate is **parity unpinned** (SURVEY §8c).  The predicate stated here, in this operation order, is what the
CUDA kernel must reproduce bit for bit:

    d = p − translation                      (f32)
    l = R · d,  l_r = (R_r0·d0 + R_r1·d1) + R_r2·d2            (f32, no contraction)
    ball      (l0² + l1²) + l2² <= r·r
    cuboid    |l_a| <= half_extent_a for every axis
    cylinder  |l1| <= half_height  and  l0² + l2² <= r·r       (axis = local Y)
"""
from __future__ import annotations

import numpy as np

F = np.float32


def world_points(m: np.ndarray, nx: int, ny: int, z_begin: int, z_end: int):
    """f64 row-by-row affine map of every lattice point of planes [z_begin, z_end), cast to f32.  Returns three
    broadcastable [nz, ny, nx] arrays."""
    x = np.arange(nx, dtype=np.float64)[None, None, :]
    y = np.arange(ny, dtype=np.float64)[None, :, None]
    z = np.arange(z_begin, z_end, dtype=np.float64)[:, None, None]
    out = []
    for r in range(3):
        w = ((m[r, 0] * x + m[r, 1] * y) + m[r, 2] * z) + m[r, 3]
        out.append(np.broadcast_to(w, (z_end - z_begin, ny, nx)).astype(F))
    return out


def contains(kind: int, params, rot: np.ndarray, translation, px, py, pz) -> np.ndarray:
    rot = np.asarray(rot, dtype=F).reshape(3, 3)
    t = np.asarray(translation, dtype=F)
    d = (px - t[0], py - t[1], pz - t[2])
    l = [(rot[r, 0] * d[0] + rot[r, 1] * d[1]) + rot[r, 2] * d[2] for r in range(3)]
    p = [F(v) for v in params]
    if kind == 0:
        return (l[0] * l[0] + l[1] * l[1]) + l[2] * l[2] <= p[0] * p[0]
    if kind == 1:
        return (np.abs(l[0]) <= p[0]) & (np.abs(l[1]) <= p[1]) & (np.abs(l[2]) <= p[2])
    if kind == 2:
        return (np.abs(l[1]) <= p[0]) & (l[0] * l[0] + l[2] * l[2] <= p[1] * p[1])
    raise ValueError(kind)


def rasterize(lattice, solver_to_world: np.ndarray, shapes, default_label: int = 0, z_begin: int = 0, z_end=None) -> np.ndarray:
    """shapes: [(kind, params, world_to_local_rot3x3_f32, translation, label)] in priority order.  Returns u32 labels, flat,
    x fastest.  First match wins (shapes are applied last-to-first so earlier ones overwrite later ones)."""
    nx, ny, nz = lattice
    z_end = nz if z_end is None else z_end
    m = np.asarray(solver_to_world, dtype=np.float64).reshape(4, 4)
    px, py, pz = world_points(m, nx, ny, z_begin, z_end)
    out = np.full((z_end - z_begin, ny, nx), default_label, dtype=np.uint32)
    for kind, params, rot, translation, label in reversed(list(shapes)):
        out[contains(kind, params, rot, translation, px, py, pz)] = label
    return out.reshape(-1)


def rasterize_scene_domain(scene, z_begin: int = 0, z_end=None) -> np.ndarray:
    """Convenience for tests: takes a fdtd_b200.scene.SceneDomain (data only — nothing of the product is executed)."""
    shapes = [(s.kind, list(s.params) + [0.0] * (3 - len(s.params)), s.world_to_local_rotation(), s.translation, lab)
              for s, lab in zip(scene.shapes, scene.labels)]
    return rasterize(scene.transformations.lattice_size, scene.transformations.transform_from_solver_to_world, shapes, 0, z_begin, z_end)
