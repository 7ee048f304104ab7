//! `Field::field → FieldView::{at, iter}` (cem-solver/src/lib.rs:97-136) over `fdtd_b200_read_field`.
//!
//! Like the reference's wgpu backend (cem-solver/src/fdtd/wgpu/mod.rs:451-578) the view owns a host copy of the box, read
//! back once when the view is created, and widens f32 → f64 on access (wgpu/mod.rs:522, 565).  The box is the reference's
//! `normalize_point_bounds` (util.rs:140-162) clipped to the lattice; iteration is x-fastest (util.rs:174-199).  The data
//! is the FRESH field (buffer `tick % 2` of the reference, cpu/mod.rs:538).
use std::ops::{Range, RangeBounds};

use cem_solver::{Field, FieldComponent, FieldView};
use nalgebra::{Point3, Vector3};

use crate::{FdtdB200SolverInstance, FdtdB200SolverState, error::last_error, ffi};

pub(crate) fn ffi_component(component: FieldComponent) -> std::os::raw::c_int {
    match component {
        FieldComponent::E => ffi::FDTD_B200_FIELD_E,
        FieldComponent::H => ffi::FDTD_B200_FIELD_H,
    }
}

impl Field<Point3<usize>> for FdtdB200SolverInstance {
    type View<'a>
        = B200FieldView
    where
        Self: 'a;

    fn field<'a, R>(
        &'a self,
        state: &'a FdtdB200SolverState,
        range: R,
        field_component: FieldComponent,
    ) -> B200FieldView
    where
        R: RangeBounds<Point3<usize>>,
    {
        let range = self.lattice().normalize_bounds(&range);
        let extent = range.end - range.start;
        let n_cells = extent.x * extent.y * extent.z;
        let mut data = vec![0.0f32; 3 * n_cells];
        if n_cells > 0 {
            let lo = [range.start.x as u64, range.start.y as u64, range.start.z as u64];
            let hi = [range.end.x as u64, range.end.y as u64, range.end.z as u64];
            // blocks until the box is on the host, like the staging-buffer map of the wgpu backend (typed.rs:407-471)
            let rc = unsafe {
                ffi::fdtd_b200_read_field(
                    state.h,
                    ffi_component(field_component),
                    lo.as_ptr(),
                    hi.as_ptr(),
                    data.as_mut_ptr(),
                )
            };
            assert_eq!(rc, ffi::FDTD_B200_OK, "fdtd_b200_read_field failed: {}", last_error(state.h));
        }
        B200FieldView { range, extent, data }
    }
}

/// Host copy of a box of E or H, xyz interleaved, x fastest.
#[derive(Clone, Debug)]
pub struct B200FieldView {
    range: Range<Point3<usize>>,
    extent: Vector3<usize>,
    data: Vec<f32>,
}

impl B200FieldView {
    fn value(&self, offset: usize) -> Vector3<f64> {
        Vector3::new(
            self.data[3 * offset] as f64,
            self.data[3 * offset + 1] as f64,
            self.data[3 * offset + 2] as f64,
        )
    }

    pub fn range(&self) -> &Range<Point3<usize>> {
        &self.range
    }
}

impl FieldView<Point3<usize>> for B200FieldView {
    type Iter<'a>
        = B200FieldIter<'a>
    where
        Self: 'a;

    fn at(&self, point: &Point3<usize>) -> Option<Vector3<f64>> {
        let inside = (0..3).all(|a| point[a] >= self.range.start[a] && point[a] < self.range.end[a]);
        inside.then(|| {
            let p = point - self.range.start;
            self.value(p.x + self.extent.x * (p.y + self.extent.y * p.z))
        })
    }

    fn iter<'a>(&'a self) -> B200FieldIter<'a> {
        B200FieldIter { view: self, next: 0 }
    }
}

/// x-fastest walk over the box, yielding lattice points (not box-relative ones) like `CpuFieldView::iter`
/// (cem-solver/src/fdtd/cpu/mod.rs:550-595).
#[derive(Clone, Debug)]
pub struct B200FieldIter<'a> {
    view: &'a B200FieldView,
    next: usize,
}

impl Iterator for B200FieldIter<'_> {
    type Item = (Point3<usize>, Vector3<f64>);

    fn next(&mut self) -> Option<Self::Item> {
        let e = &self.view.extent;
        let n = e.x * e.y * e.z;
        if self.next >= n {
            return None;
        }
        let i = self.next;
        self.next += 1;
        let plane = e.x * e.y;
        let offset = Vector3::new(i % e.x, (i % plane) / e.x, i / plane);
        Some((self.view.range.start + offset, self.view.value(i)))
    }

    fn size_hint(&self) -> (usize, Option<usize>) {
        let e = &self.view.extent;
        let left = e.x * e.y * e.z - self.next;
        (left, Some(left))
    }
}

impl ExactSizeIterator for B200FieldIter<'_> {}
