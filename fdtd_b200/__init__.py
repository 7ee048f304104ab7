"""fdtd_b200 — B200-native (sm_100a) replacement for cem-solver's FDTD time-stepping path.

Layout:  csrc/ (CUDA kernels + the C ABI of include/fdtd_b200.h), _lib.py (ctypes binding),
solver.py (host-side mirror of the reference's solver traits), driver.py (headless solver loop),
configs.py (the BASELINE.json workloads), distributed.py (z-slab decomposition + halo exchange).
"""
from ._lib import (COLORMAP_APP_Z, COLORMAP_CPU, COLORMAP_LINEAR, EXACT_ARITHMETIC, FORCE_GENERIC_KERNELS, FUSED_STEP, IN_PLACE_STEP,
                   Fdtd_b200Error, lib)
from .solver import (B200FieldView, B200Projection, B200ProjectionPass, ContinousWave, DomainDescription,
                     FdtdB200Backend, FdtdB200SolverInstance, FdtdB200SolverState, FdtdB200UpdatePass,
                     FdtdSolverConfig, FieldComponent, GaussianPulse, GradedPml, ImageTarget, Material, MaterialLattice,
                     PhysicalConstants, PmlCoefficients, PmlProfiles, ProjectionParameters, Resolution, SourceValues, WithAmplitudes,
                     normalize_point_bounds)

__all__ = [n for n in dir() if not n.startswith("_")]
