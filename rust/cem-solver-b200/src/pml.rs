//! `DomainDescription::pml(point) -> Option<PmlCoefficients>` (cem-solver/src/lib.rs:60-67) folded into what the C ABI takes:
//! per axis a profile (b[i], c[i]) over the coordinate along that axis (`fdtd_b200_set_pml`).
//!
//! The reference collects the per-cell answers (cpu/mod.rs:242-255) and never uses them (`psi = zeros`,
//! cpu/mod.rs:419-425 "todo: use pml"); the B200 engine runs the CPML recursion, for separable PMLs.  The application's
//! answers are separable by construction (axis-aligned slabs, cem-app/src/solver/runner.rs:633-675) but carry the sign of
//! `-normal` (runner.rs:667), so b and c come out negative on three faces: magnitudes are used (b >= 0, c <= 0), as the
//! Python mirror does (fdtd_b200/solver.py `PmlProfiles.from_cells`).
use cem_solver::fdtd::pml::PmlCoefficients;
use nalgebra::{Point3, Vector3};

use crate::error::B200Error;

/// Per-axis profiles; `None` where an axis has no PML coordinate.
pub type PmlProfiles = [Option<AxisProfile>; 3];

#[derive(Clone, Debug, PartialEq)]
pub struct AxisProfile {
    pub b: Vec<f64>,
    pub c: Vec<f64>,
}

#[derive(Clone, Debug)]
pub struct PmlProfileBuilder {
    b: [Vec<f64>; 3],
    c: [Vec<f64>; 3],
    conflict: Option<String>,
}

impl PmlProfileBuilder {
    pub fn new(size: &Vector3<usize>) -> Self {
        Self {
            b: [vec![0.0; size.x], vec![0.0; size.y], vec![0.0; size.z]],
            c: [vec![0.0; size.x], vec![0.0; size.y], vec![0.0; size.z]],
            conflict: None,
        }
    }

    /// One `Some(coefficients)` answer of `DomainDescription::pml`.
    pub fn add(&mut self, point: &Point3<usize>, coefficients: &PmlCoefficients) {
        for axis in 0..3 {
            let b = coefficients.b[axis].abs();
            let c = -coefficients.c[axis].abs();
            if c == 0.0 {
                continue;
            }
            let i = point[axis];
            if self.c[axis][i] != 0.0 && (self.b[axis][i] != b || self.c[axis][i] != c) {
                self.conflict.get_or_insert_with(|| {
                    format!("axis {axis} coordinate {i} was given two different coefficient sets (second at {point:?})")
                });
                continue;
            }
            self.b[axis][i] = b;
            self.c[axis][i] = c;
        }
    }

    pub fn finish(self) -> Result<PmlProfiles, B200Error> {
        if let Some(message) = self.conflict {
            return Err(B200Error::PmlNotSeparable(message));
        }
        let [b0, b1, b2] = self.b;
        let [c0, c1, c2] = self.c;
        let axis = |b: Vec<f64>, c: Vec<f64>| c.iter().any(|&v| v != 0.0).then_some(AxisProfile { b, c });
        Ok([axis(b0, c0), axis(b1, c1), axis(b2, c2)])
    }
}
