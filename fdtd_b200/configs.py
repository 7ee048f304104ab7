"""The BASELINE.json workloads made concrete: lattice, palette + material-id lattice,
sources, probes, step count.  Pure host-side numpy; the rasterisation rule is the reference's
"first entity whose collider contains the point wins, else the default material"
(cem-app/src/solver/runner.rs:616-631) on cell coordinates (dx = 1, solver == world units).
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import Optional, Sequence

import numpy as np

from .solver import (ContinousWave, FdtdSolverConfig, GaussianPulse, Material, MaterialLattice, PhysicalConstants,
                     Resolution, WithAmplitudes)


@dataclass
class Workload:
    name: str
    lattice: tuple                      # (nx, ny, nz)
    resolution: Resolution
    constants: PhysicalConstants
    palette: list                       # list[Material], canonical (first-occurrence) order
    ids: Optional[np.ndarray]           # uint32 [nz*ny*nx] or None (all palette[0])
    sources: list                       # [(point, WithAmplitudes)]
    probes: list = field(default_factory=list)   # [point]
    steps: int = 0
    scene: Optional[object] = None      # scene.SceneDomain: the lattice is rasterised on the device (then `ids` is None)
    pml: Optional[object] = None        # solver.PmlProfiles: CPML slabs; None = no PML, as the reference executes

    def solver_config(self) -> FdtdSolverConfig:
        return FdtdSolverConfig.from_lattice(self.lattice, self.resolution, self.constants)

    def domain(self):
        if self.scene is not None:
            self.scene.pml_profiles = self.pml
            return self.scene
        return MaterialLattice(self.palette, self.ids, pml=self.pml)

    @property
    def n_cells(self) -> int:
        return self.lattice[0] * self.lattice[1] * self.lattice[2]

    def palette_rows(self) -> np.ndarray:
        return np.array([m.as_row() for m in self.palette], dtype=np.float64)

    def source_table(self, n_steps: int, t0: float = 0.0) -> np.ndarray:
        """[n_steps, n_src, 6] f64, evaluated at the accumulated pre-step time t_n (runner.rs:522-526;
        time += dt as cpu/mod.rs:513 — not tick*dt)."""
        out = np.zeros((n_steps, len(self.sources), 6), dtype=np.float64)
        t = float(t0)
        for s in range(n_steps):
            for k, (_, src) in enumerate(self.sources):
                v = src.evaluate(t)
                out[s, k, :3] = v.j
                out[s, k, 3:] = v.m
            t += self.resolution.temporal
        return out

    def source_points(self) -> list:
        return [p for p, _ in self.sources]

    def cell(self, p) -> int:
        return int(p[0]) + self.lattice[0] * (int(p[1]) + self.lattice[1] * int(p[2]))


REDUCED_RES = Resolution((1.0, 1.0, 1.0), 0.25)  # the UI's default solver config (cem-app/src/solver/ui.rs:283-286)


def canonicalize(labels: np.ndarray, palette: Sequence[Material]) -> tuple:
    """First-occurrence renumbering of a labelled lattice, in numpy (host-side scene prep)."""
    flat = labels.reshape(-1)
    uniq, first = np.unique(flat, return_index=True)
    order = uniq[np.argsort(first)]
    remap = np.zeros(int(flat.max()) + 1, dtype=np.uint32)
    remap[order] = np.arange(len(order), dtype=np.uint32)
    return remap[flat], [palette[int(k)] for k in order]


def _ball_mask_apply(labels: np.ndarray, free: np.ndarray, centre, radius: float, label: int) -> None:
    nz, ny, nx = labels.shape
    cx, cy, cz = centre
    r = int(np.ceil(radius)) + 1
    x0, x1 = max(int(cx) - r, 0), min(int(cx) + r + 1, nx)
    y0, y1 = max(int(cy) - r, 0), min(int(cy) + r + 1, ny)
    z0, z1 = max(int(cz) - r, 0), min(int(cz) + r + 1, nz)
    if x0 >= x1 or y0 >= y1 or z0 >= z1:
        return
    z, y, x = np.ogrid[z0:z1, y0:y1, x0:x1]
    inside = (x - cx) ** 2 + (y - cy) ** 2 + (z - cz) ** 2 <= radius * radius
    sub_free = free[z0:z1, y0:y1, x0:x1]
    hit = inside & sub_free
    labels[z0:z1, y0:y1, x0:x1][hit] = label
    sub_free[hit] = False


def _box_mask_apply(labels: np.ndarray, free: np.ndarray, lo, hi, label: int) -> None:
    nz, ny, nx = labels.shape
    x0, y0, z0 = (max(int(v), 0) for v in lo)
    x1, y1, z1 = min(int(hi[0]), nx), min(int(hi[1]), ny), min(int(hi[2]), nz)
    if x0 >= x1 or y0 >= y1 or z0 >= z1:
        return
    sub_free = free[z0:z1, y0:y1, x0:x1]
    labels[z0:z1, y0:y1, x0:x1][sub_free] = label
    sub_free[...] = False


def config1_vacuum64(n: int = 64, steps: int = 1000) -> Workload:
    """64³ vacuum cube, Gaussian-pulse point source (presets.rs:176-178 rescaled to dt=0.25), 1000 steps."""
    c = n // 2
    src = GaussianPulse(12.5, 2.5).with_amplitudes((0.0, 0.0, 50.0), (0.0, 0.0, 0.0))
    return Workload(f"vacuum{n}", (n, n, n), REDUCED_RES, PhysicalConstants.REDUCED, [Material.VACUUM], None,
                    [((c, c, c), src)], [], steps)


def config2_inclusions(n: int = 256, steps: int = 5000) -> Workload:
    """256³ with dielectric / lossy / magnetic inclusions, a conductor slab and 8 probes (scaled with n)."""
    s = n / 256.0
    pal = [Material.VACUUM,
           Material(relative_permittivity=3.9),                                   # presets.rs:70-73
           Material(eletrical_conductivity=10.0),                                  # composer/mod.rs:855
           Material(relative_permeability=2.0, magnetic_conductivity=0.5),
           Material(eletrical_conductivity=1.0e4)]
    labels = np.zeros((n, n, n), dtype=np.uint32)
    free = np.ones((n, n, n), dtype=bool)
    _ball_mask_apply(labels, free, (96 * s, 128 * s, 128 * s), 40 * s, 1)
    _box_mask_apply(labels, free, ((176 - 24) * s, (128 - 24) * s, (96 - 24) * s), ((176 + 24) * s, (128 + 24) * s, (96 + 24) * s), 2)
    _ball_mask_apply(labels, free, (128 * s, 64 * s, 176 * s), 24 * s, 3)
    _box_mask_apply(labels, free, (0, 0, int(200 * s)), (n, n, max(int(204 * s), int(200 * s) + 1)), 4)
    ids, pal2 = canonicalize(labels, pal)
    c = n // 2
    src = ContinousWave(0.0, 0.05 / s).with_amplitudes((0.0, 0.0, 50.0), (0.0, 0.0, 0.0))
    probes = [(int((64 * k + 32) * s), c, c) for k in range(4)] + [(c, c, int((64 * k + 32) * s)) for k in range(4)]
    return Workload(f"inclusions{n}", (n, n, n), REDUCED_RES, PhysicalConstants.REDUCED, pal2, ids, [((c, c, c), src)], probes, steps)


def absorbing_layer_labels(shape_zyx: tuple, depth: int = 16) -> np.ndarray:
    """Layer index 0 (interior) .. depth per cell: d = (depth - dist_to_nearest_face) max over axes."""
    nz, ny, nx = shape_zyx

    def axis_level(n: int) -> np.ndarray:
        i = np.arange(n)
        dist = np.minimum(i, n - 1 - i)
        return np.clip(depth - dist, 0, depth).astype(np.uint32)

    lz, ly, lx = axis_level(nz), axis_level(ny), axis_level(nx)
    return np.maximum(np.maximum(lz[:, None, None], ly[None, :, None]), lx[None, None, :])


def layer_materials(depth: int, constants: PhysicalConstants, sigma_max: float = 2.5, m: float = 4.0) -> list:
    """The only absorber the reference executes is material conductivity: graded, impedance
    matched lossy layer with the GradedPml grading σ = σmax·d^m (presets.rs:123-130, pml.rs:80-84)."""
    out = []
    for level in range(1, depth + 1):
        d = level / depth
        sigma = sigma_max * d ** m
        out.append(Material(1.0, sigma * constants.vacuum_permeability / constants.vacuum_permittivity, 1.0, sigma))
    return out


def config3_heterogeneous(n: int = 512, steps: int = 200, n_balls: int = 64, seed: int = 12345, nz: Optional[int] = None) -> Workload:
    """512³ heterogeneous lattice (64 random balls of 4 material kinds) inside a 16-cell graded lossy layer."""
    nz = n if nz is None else nz
    s = n / 512.0
    depth = max(int(round(16 * s)), 2)
    kinds = [Material(relative_permittivity=3.9), Material(relative_permittivity=2.2, eletrical_conductivity=0.05),
             Material(relative_permeability=2.0, magnetic_conductivity=0.5), Material(eletrical_conductivity=10.0)]
    layer = layer_materials(depth, PhysicalConstants.REDUCED)
    pal = [Material.VACUUM] + layer + kinds
    labels = absorbing_layer_labels((nz, n, n), depth)            # 0 interior, 1..depth layer
    free = labels == 0                                            # inclusions excluded from the layer
    rng = np.random.default_rng(seed)
    radii = rng.uniform(8 * s, 40 * s, n_balls)
    centres = rng.uniform(56 * s, 456 * s, (n_balls, 3))
    if nz != n:
        centres[:, 2] *= nz / n
    for b in range(n_balls):
        _ball_mask_apply(labels, free, centres[b], radii[b], 1 + depth + (b % 4))
    ids, pal2 = canonicalize(labels, pal)
    src = GaussianPulse(12.5, 2.5).with_amplitudes((0.0, 0.0, 50.0), (0.0, 0.0, 0.0))
    c = (n // 2, n // 2, nz // 2)
    return Workload(f"heterogeneous{n}x{n}x{nz}", (n, n, nz), REDUCED_RES, PhysicalConstants.REDUCED, pal2, ids, [(c, src)], [], steps)


def config4_slab_lattice(nx: int, ny: int, nz: int, steps: int = 100) -> Workload:
    """Multi-GPU weak/strong scaling lattice: vacuum inside the config-3 lossy layer.  The id lattice is
    produced per z-slab by `slab_ids` so no rank materialises the whole 2048³ index."""
    depth = 16
    pal = [Material.VACUUM] + layer_materials(depth, PhysicalConstants.REDUCED)
    src = GaussianPulse(12.5, 2.5).with_amplitudes((0.0, 0.0, 50.0), (0.0, 0.0, 0.0))
    w = Workload(f"layer{nx}x{ny}x{nz}", (nx, ny, nz), REDUCED_RES, PhysicalConstants.REDUCED, pal, None,
                 [((nx // 2, ny // 2, nz // 2), src)], [], steps)
    return w


def slab_layer_ids(lattice: tuple, z_begin: int, z_end: int, depth: int = 16) -> np.ndarray:
    """Material ids of planes [z_begin, z_end) of the config-4 lattice.  Labels equal layer level (0 = vacuum)
    = the palette order of `config4_slab_lattice` (palette[k] = level k; deliberately not renumbered by first
    occurrence — the C ABI accepts any labelling and each rank only ever sees its own slab)."""
    nx, ny, nz = lattice

    def axis_level(n: int, idx: np.ndarray) -> np.ndarray:
        dist = np.minimum(idx, n - 1 - idx)
        return np.clip(depth - dist, 0, depth).astype(np.uint32)

    lz = axis_level(nz, np.arange(z_begin, z_end))
    ly = axis_level(ny, np.arange(ny))
    lx = axis_level(nx, np.arange(nx))
    lab = np.maximum(np.maximum(lz[:, None, None], ly[None, :, None]), lx[None, None, :])
    return np.ascontiguousarray(lab.reshape(-1), dtype=np.uint32)


def heterogeneous_palette(depth: int = 16) -> list:
    """palette[0] vacuum, [1..depth] absorbing-layer levels, then the four inclusion kinds (labels of
    `heterogeneous_slab_labels`; not renumbered — every rank must share one table)."""
    kinds = [Material(relative_permittivity=3.9), Material(relative_permittivity=2.2, eletrical_conductivity=0.05),
             Material(relative_permeability=2.0, magnetic_conductivity=0.5), Material(eletrical_conductivity=10.0)]
    return [Material.VACUUM] + layer_materials(depth, PhysicalConstants.REDUCED) + kinds


def heterogeneous_slab_labels(lattice: tuple, z_begin: int, z_end: int, n_balls: int = 64, seed: int = 12345, depth: int = 16) -> np.ndarray:
    """Planes [z_begin, z_end) of a config-3 style lattice of any shape (balls scaled with the lattice), so that
    each rank of a slab-decomposed run rasterises only what it owns.  Returns uint32 labels into
    `heterogeneous_palette(depth)`, flat, x fastest."""
    nx, ny, nz = lattice

    def axis_level(n: int, idx: np.ndarray) -> np.ndarray:
        dist = np.minimum(idx, n - 1 - idx)
        return np.clip(depth - dist, 0, depth).astype(np.uint32)

    lz = axis_level(nz, np.arange(z_begin, z_end))
    labels = np.maximum(np.maximum(lz[:, None, None], axis_level(ny, np.arange(ny))[None, :, None]),
                        axis_level(nx, np.arange(nx))[None, None, :])
    labels = np.ascontiguousarray(labels)
    free = labels == 0
    rng = np.random.default_rng(seed)
    scale = min(nx, ny, nz) / 512.0
    radii = rng.uniform(8 * scale, 40 * scale, n_balls)
    centres = rng.uniform(56 / 512.0, 456 / 512.0, (n_balls, 3)) * np.array([nx, ny, nz])
    for b in range(n_balls):
        c = centres[b].copy()
        c[2] -= z_begin  # slab-local z
        _ball_mask_apply(labels, free, c, radii[b], 1 + depth + (b % 4))
    return labels.reshape(-1)


# ---- config 5: dipole antenna from a NEC deck -------------------------------------------------------------
NEC_DIPOLE_DECK = """CM half-wave dipole
CE
GW 1 21 0 0 -0.25 0 0 0.25 0.005
GE 0
EX 0 1 11 0 1 0
RP 0 19 1 1000 0 0 5 0
EN
"""


def config5_nec_dipole(n: int = 256, steps: int = 2000, deck: str = NEC_DIPOLE_DECK) -> Workload:
    """The deck goes through the NEC parser/interpreter (nec.py) and becomes 21 Y-aligned cylinders exactly as the
    application lays them out (composer/file_formats/nec.rs:47-70 — strung along +Y from the wire's first end whatever the
    wire's direction); wire segments are σ = 1e4 conductor except the centre segment, which stays vacuum and holds the feed:
    a Gaussian j_y at its centre cell.  The lattice is centred on the wire, dx = wire length / (63·n/256) so the wire spans
    a quarter of it; "far-field" probes = 16 cells on a circle of radius 100·n/256 in the xz-plane + 2 on the wire axis (the
    reference has no near-to-far-field transform and reads no EX/RP cards, nec-file/src/parser.rs:91-100)."""
    from .nec import NecFile, wire_cylinders
    from .scene import CoordinateTransformations, SceneDomain, Shape

    cyl = wire_cylinders(NecFile.from_text(deck))
    length = float(sum(2.0 * float(c.half_height) for c in cyl))
    s = n / 256.0
    dx = length / (63.0 * s)
    res = Resolution((dx, dx, dx), 0.25 * dx)
    mid = cyl[len(cyl) // 2]
    c = n // 2
    mins = tuple(float(mid.centre[a]) - c * dx for a in range(3))
    tf = CoordinateTransformations.for_fdtd(res, (n, n, n), aabb_mins=mins)
    wire = Material(eletrical_conductivity=1.0e4)
    shapes = [Shape.cylinder(float(k.half_height), float(k.radius), Material.VACUUM if k is mid else wire,
                             tuple(float(v) for v in k.centre)) for k in cyl]
    scene = SceneDomain(shapes, tf, Material.VACUUM)
    src = GaussianPulse(50 * res.temporal, 10 * res.temporal).with_amplitudes((0.0, 50.0, 0.0), (0.0, 0.0, 0.0))
    r = int(round(100 * s))
    probes = [(c + int(round(r * np.cos(2 * np.pi * k / 16))), c, c + int(round(r * np.sin(2 * np.pi * k / 16)))) for k in range(16)]
    probes += [(c, c + r, c), (c, c - r, c)]
    return Workload(f"nec_dipole{n}", (n, n, n), res, PhysicalConstants.REDUCED, scene.palette, None, [((c, c, c), src)], probes, steps,
                    scene=scene)
