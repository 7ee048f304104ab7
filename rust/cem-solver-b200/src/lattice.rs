//! Lattice indexing.  cem-solver keeps `Strider` and `normalize_point_bounds` private (`mod strider; mod util;` in
//! cem-solver/src/fdtd/mod.rs:4-5), so the few lines this crate needs are restated: linear index = x + nx·(y + ny·z)
//! (cem-solver/src/fdtd/strider.rs:34-53, 147-154), bounds normalisation as cem-solver/src/fdtd/util.rs:140-162.
use std::ops::{Bound, Range, RangeBounds};

use nalgebra::{Point3, Vector3};

#[derive(Clone, Copy, Debug, PartialEq, Eq)]
pub struct Lattice {
    pub size: Vector3<usize>,
}

impl Lattice {
    pub fn new(size: Vector3<usize>) -> Self {
        Self { size }
    }

    pub fn len(&self) -> usize {
        self.size.x * self.size.y * self.size.z
    }

    pub fn is_empty(&self) -> bool {
        self.len() == 0
    }

    pub fn is_inside(&self, point: &Point3<usize>) -> bool {
        point.x < self.size.x && point.y < self.size.y && point.z < self.size.z
    }

    /// `Strider::index` (strider.rs:51-53): `None` outside the lattice.
    pub fn index(&self, point: &Point3<usize>) -> Option<usize> {
        self.is_inside(point)
            .then(|| point.x + self.size.x * (point.y + self.size.y * point.z))
    }

    /// `Strider::point_unchecked` (strider.rs:34-41).
    pub fn point_unchecked(&self, index: usize) -> Point3<usize> {
        let plane = self.size.x * self.size.y;
        let z = index / plane;
        let rest = index % plane;
        Point3::new(rest % self.size.x, rest / self.size.x, z)
    }

    /// `normalize_point_bounds` (util.rs:140-162) followed by the clip to the lattice that the C ABI applies
    /// (`fdtd_b200_read_field`: "half-open box clipped to the lattice").
    pub fn normalize_bounds(&self, range: &impl RangeBounds<Point3<usize>>) -> Range<Point3<usize>> {
        let start = match range.start_bound() {
            Bound::Included(start) => *start,
            Bound::Excluded(start) => start + Vector3::repeat(1),
            Bound::Unbounded => Point3::origin(),
        };
        let end = match range.end_bound() {
            Bound::Included(end) => end + Vector3::repeat(1),
            Bound::Excluded(end) => *end,
            Bound::Unbounded => self.size.into(),
        };
        let start: Point3<usize> = start.coords.zip_map(&self.size, |x, n| x.min(n)).into();
        let end: Point3<usize> = Point3::from(
            start
                .coords
                .zip_map(&end.coords, |x0, x1| x0.max(x1))
                .zip_map(&self.size, |x, n| x.min(n)),
        );
        Range { start, end }
    }
}

#[cfg(test)]
mod tests {
    use super::*;

    #[test]
    fn index_is_x_fastest() {
        let l = Lattice::new(Vector3::new(4, 3, 2));
        assert_eq!(l.index(&Point3::new(1, 2, 1)), Some(1 + 4 * (2 + 3)));
        assert_eq!(l.index(&Point3::new(4, 0, 0)), None);
        assert_eq!(l.point_unchecked(21), Point3::new(1, 2, 1));
    }

    #[test]
    fn bounds_follow_the_reference() {
        let l = Lattice::new(Vector3::new(4, 3, 2));
        let r = l.normalize_bounds(&(Point3::new(1, 1, 0)..=Point3::new(2, 2, 1)));
        assert_eq!(r, Point3::new(1, 1, 0)..Point3::new(3, 3, 2));
        let r = l.normalize_bounds(&(..));
        assert_eq!(r, Point3::origin()..Point3::new(4, 3, 2));
        // end < start collapses to an empty box (util.rs:157-160)
        let r = l.normalize_bounds(&(Point3::new(2, 2, 1)..Point3::new(1, 1, 0)));
        assert_eq!(r.start, r.end);
    }
}
