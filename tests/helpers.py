"""Shared helpers for the parity tests: run the same workload through the oracle and the CUDA path."""
from __future__ import annotations

import numpy as np


def rel_l2(a: np.ndarray, b: np.ndarray) -> float:
    a = np.asarray(a, dtype=np.float64)
    b = np.asarray(b, dtype=np.float64)
    d = np.linalg.norm((a - b).ravel())
    n = np.linalg.norm(b.ravel())
    return float(d / n) if n > 0 else float(d)


def oracle_run(oracle, w, n_steps: int, precision: int, init=None, probes=None, fast: bool = False):
    """Returns (solver, probe_series).  init = (E0, H0) arrays [nz,ny,nx,3] or None."""
    s = oracle.RefSolver(w.lattice, w.resolution.spatial, w.resolution.temporal, precision=precision, fast=fast)
    s.set_materials(w.palette_rows(), w.ids, w.constants.vacuum_permittivity, w.constants.vacuum_permeability)
    if getattr(w, "pml", None) is not None and w.pml.any():
        n = w.lattice
        ax = [a if a is not None else (np.zeros(n[k]),) * 4 for k, a in enumerate(w.pml.axes)]
        s.set_pml_profiles([a[0] for a in ax], [a[1] for a in ax], [a[2] for a in ax], [a[3] for a in ax])
    if init is not None:
        s.write_field(oracle.E, init[0])
        s.write_field(oracle.H, init[1])
    jm = w.source_table(n_steps) if w.sources else None
    cells = [w.cell(p) for p in probes] if probes else None
    series, _ = s.run(n_steps, w.source_points() if w.sources else None, jm, cells)
    return s, series


def b200_state(b200, w, exact: bool = False, generic: bool = False, z_slab=None, device: int = 0, fused: bool = False):
    backend = b200.FdtdB200Backend(device=device, exact_arithmetic=exact, force_generic_kernels=generic, z_slab=z_slab, fused_step=fused)
    inst = backend.create_instance(w.solver_config(), w.domain())
    return inst, inst.create_state()


def b200_run(b200, w, n_steps: int, exact: bool = False, generic: bool = False, init=None, probes=None, fused: bool = False):
    inst, st = b200_state(b200, w, exact, generic, fused=fused)
    if init is not None:
        st.write_field(b200.FieldComponent.E, init[0])
        st.write_field(b200.FieldComponent.H, init[1])
    if probes:
        st.register_probes(probes)
    if w.sources:
        st.set_source_table(w.source_points(), w.source_table(n_steps))
    st.step(n_steps)
    st.sync()
    series = st.read_probes(0, n_steps) if probes else None
    return inst, st, series


def random_fields(w, seed: int = 0):
    rng = np.random.default_rng(seed)
    nx, ny, nz = w.lattice
    e = rng.standard_normal((nz, ny, nx, 3)).astype(np.float32)
    h = rng.standard_normal((nz, ny, nx, 3)).astype(np.float32)
    return e, h
