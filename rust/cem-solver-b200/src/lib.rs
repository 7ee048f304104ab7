//! `FdtdB200Backend`: cem-solver's solver traits (cem-solver/src/lib.rs:28-142) over libfdtd_b200.so, the sm_100a CUDA
//! engine of the fdtd_b200 repository (C ABI: include/fdtd_b200.h, declared in [`ffi`]).
//!
//! Mapping (reference → here):
//! * `SolverBackend::create_instance` (lib.rs:28-45): the per-cell `material(point)` / `pml(point)` loop of
//!   cpu/mod.rs:244-264 runs unchanged, in linear-index order; `fdtd_b200_index_materials` compresses it into first-occurrence
//!   material ids + a palette (README.md:25-28 TODO), `pml` answers are folded into per-axis profiles ([`pml`]).
//! * `SolverBackend::memory_required` (lib.rs:41-44) = `fdtd_b200_memory_required` in the step mode the backend will use.
//!   The runner compares it with `memory_limit` before construction (cem-app/src/solver/runner.rs:304-330); a backend built
//!   with [`FdtdB200Backend::with_memory_limit`] takes the fused single-pass step if its two field sets fit the limit, else the
//!   in-place two-kernel step (`fdtd_b200_choose_step_mode`).  No flag = the library's default = the fused step.
//! * `SolverInstance::create_state` (lib.rs:70-79) = `fdtd_b200_create` (one handle = the reference's Instance + State).
//! * `begin_update` / `UpdatePassForcing::set_forcing` / `UpdatePass::finish` (lib.rs:81-90; cpu/mod.rs:370-515) =
//!   collect `fdtd_b200_source`s → `fdtd_b200_set_sources` → `fdtd_b200_step(h, 1)`.
//! * `Time` (lib.rs:92-95) = `fdtd_b200_tick` / `fdtd_b200_time`.
//! * `Field` / `FieldView` (lib.rs:97-136): [`field`]; projections (project.rs:75-133): [`project`].
//!
//! Threading: the handle is single-caller but may move between threads (it calls `cudaSetDevice` on entry), which is what
//! the runner needs: constructed on the UI thread, driven by the "solver" thread (runner.rs:340-352, 471).
pub mod error;
pub mod ffi;
pub mod field;
pub mod lattice;
pub mod pml;
pub mod project;

use std::os::raw::c_int;

use cem_solver::{
    DomainDescription, SolverBackend, SolverInstance, Time, UpdatePass, UpdatePassForcing,
    fdtd::FdtdSolverConfig,
    material::{Material, PhysicalConstants},
    source::SourceValues,
};
use nalgebra::Point3;

pub use crate::error::B200Error;
use crate::{
    error::{check, check_abi, check_handle, global_error, last_error},
    lattice::Lattice,
    pml::{PmlProfileBuilder, PmlProfiles},
};

/// Largest palette the engine indexes (16-bit material ids).
const PALETTE_CAP: usize = 65536;

/// The CUDA backend.  `Default` = device 0, fast arithmetic, the library's default step mode (fused single pass).
#[derive(Clone, Copy, Debug, Default, PartialEq, Eq)]
pub struct FdtdB200Backend {
    /// CUDA device ordinal.
    pub device: i32,
    /// Reproduce update.wgsl:89,127 operation by operation (bit-identical to an fp32 evaluation of the reference's formulas).
    pub exact_arithmetic: bool,
    /// `SolverConfigCommon::memory_limit` (cem-app/src/solver/config.rs:27-45): pick the step mode that fits.
    pub memory_limit: Option<usize>,
    /// Force the in-place two-kernel step (24 B/cell of field memory instead of 48).
    pub in_place_step: bool,
}

impl FdtdB200Backend {
    pub fn new(device: i32) -> Self {
        Self {
            device,
            ..Self::default()
        }
    }

    pub fn with_memory_limit(mut self, memory_limit: Option<usize>) -> Self {
        self.memory_limit = memory_limit;
        self
    }

    fn base_flags(&self) -> u32 {
        (if self.exact_arithmetic { ffi::FDTD_B200_EXACT_ARITHMETIC } else { 0 })
            | (if self.in_place_step { ffi::FDTD_B200_IN_PLACE_STEP } else { 0 })
    }

    /// `FdtdSolverConfig` → `fdtd_b200_config` with the step mode resolved against `memory_limit`.
    fn ffi_config(&self, config: &FdtdSolverConfig, index_bytes: u32) -> (ffi::fdtd_b200_config, usize) {
        let size = config.size(); // ceil(size / dx).max(1), fdtd/mod.rs:24-29
        let mut cfg = ffi::fdtd_b200_config {
            nx: size.x as u64,
            ny: size.y as u64,
            nz: size.z as u64,
            dx: [
                config.resolution.spatial.x,
                config.resolution.spatial.y,
                config.resolution.spatial.z,
            ],
            dt: config.resolution.temporal,
            device: self.device,
            flags: self.base_flags(),
            z_begin: 0,
            z_end: 0,
            probe_ring_steps: 0,
            reserved: 0,
        };
        let mut flags = cfg.flags;
        let mut required = 0u64;
        if !self.in_place_step {
            // fused if it fits the limit, else in place; on ENOMEM `required` is the smaller of the two and the runner's own
            // check against memory_limit (runner.rs:319-330) reports it
            let _ = unsafe {
                ffi::fdtd_b200_choose_step_mode(
                    &cfg,
                    index_bytes,
                    self.memory_limit.unwrap_or(0) as u64,
                    &mut flags,
                    &mut required,
                )
            };
            cfg.flags = flags;
        }
        else {
            required = unsafe { ffi::fdtd_b200_memory_required(&cfg, index_bytes) } as u64;
        }
        (cfg, required as usize)
    }
}

fn ffi_material(m: &Material) -> ffi::fdtd_b200_material {
    ffi::fdtd_b200_material {
        relative_permeability: m.relative_permeability,
        magnetic_conductivity: m.magnetic_conductivity,
        relative_permittivity: m.relative_permittivity,
        eletrical_conductivity: m.eletrical_conductivity,
    }
}

fn ffi_constants(c: &PhysicalConstants) -> ffi::fdtd_b200_physical_constants {
    ffi::fdtd_b200_physical_constants {
        vacuum_permittivity: c.vacuum_permittivity,
        vacuum_permeability: c.vacuum_permeability,
    }
}

impl SolverBackend<FdtdSolverConfig, Point3<usize>> for FdtdB200Backend {
    type Instance = FdtdB200SolverInstance;
    type Error = B200Error;

    fn create_instance<D>(
        &self,
        config: &FdtdSolverConfig,
        mut domain_description: D,
    ) -> Result<Self::Instance, Self::Error>
    where
        D: DomainDescription<Point3<usize>>,
    {
        check_abi()?;
        let lattice = Lattice::new(config.size());

        // the reference's loop, unchanged: pml(point) then material(point) for every linear index in order (cpu/mod.rs:244-264)
        let mut pml = PmlProfileBuilder::new(&lattice.size);
        let mut per_cell = Vec::with_capacity(lattice.len());
        for index in 0..lattice.len() {
            let point = lattice.point_unchecked(index);
            if let Some(coefficients) = domain_description.pml(&point) {
                pml.add(&point, &coefficients);
            }
            per_cell.push(ffi_material(&domain_description.material(&point)));
        }
        let pml = pml.finish()?;

        // per-cell materials → first-occurrence ids + palette (bit-exact: key = bit pattern of the four f64)
        let mut ids = vec![0u32; per_cell.len()];
        let mut palette = vec![ffi::fdtd_b200_material::default(); PALETTE_CAP];
        let mut n_palette = 0u32;
        check(unsafe {
            ffi::fdtd_b200_index_materials(
                per_cell.as_ptr(),
                per_cell.len() as u64,
                ids.as_mut_ptr(),
                palette.as_mut_ptr(),
                PALETTE_CAP as u32,
                &mut n_palette,
            )
        })?;
        palette.truncate(n_palette as usize);
        drop(per_cell);

        let index_bytes = if palette.len() + 1 <= 256 { 1 } else { 2 }; // + the pad row of padded lattices
        let (cfg, _) = self.ffi_config(config, index_bytes);
        tracing::debug!(size = ?lattice.size, materials = palette.len(), flags = cfg.flags, "creating fdtd-b200 instance");
        Ok(FdtdB200SolverInstance {
            cfg,
            constants: ffi_constants(&config.physical_constants),
            palette,
            ids,
            lattice,
            pml,
        })
    }

    fn memory_required(&self, config: &FdtdSolverConfig) -> Option<usize> {
        let (_, required) = self.ffi_config(config, 1);
        (required > 0).then_some(required)
    }
}

/// `SolverInstance`: the immutable domain (lattice, palette, material ids, PML profiles).  Plain host data, hence `Send`.
#[derive(Clone, Debug)]
pub struct FdtdB200SolverInstance {
    cfg: ffi::fdtd_b200_config,
    constants: ffi::fdtd_b200_physical_constants,
    palette: Vec<ffi::fdtd_b200_material>,
    ids: Vec<u32>,
    lattice: Lattice,
    pml: PmlProfiles,
}

impl FdtdB200SolverInstance {
    pub fn lattice(&self) -> &Lattice {
        &self.lattice
    }
}

/// `State`: owns the device-side handle (fields, index, table, sources, probes, tick, time).
#[derive(Debug)]
pub struct FdtdB200SolverState {
    pub(crate) h: ffi::fdtd_b200_handle,
}

// SAFETY: the handle is an opaque pointer to library-owned state that is only touched through `&self`/`&mut self` calls
// made by one thread at a time; the library selects its CUDA device on every entry, so the thread that created the
// handle need not be the thread that uses it (include/fdtd_b200.h, "A handle is single-caller ... may be created on one
// thread and used on another").  The runner requires `State: Send + 'static` (cem-app/src/solver/runner.rs:247-259, 443-450).
unsafe impl Send for FdtdB200SolverState {}

impl Drop for FdtdB200SolverState {
    fn drop(&mut self) {
        // SAFETY: `h` came from fdtd_b200_create and is destroyed exactly once.
        unsafe { ffi::fdtd_b200_destroy(self.h) }
    }
}

impl FdtdB200SolverState {
    /// Which kernels `finish` launches on this state and why (`fdtd_b200_step_path`) — there are no silent fallbacks.
    pub fn step_path(&self) -> String {
        let mut buf = vec![0u8; 512];
        let n = unsafe { ffi::fdtd_b200_step_path(self.h, buf.as_mut_ptr().cast(), buf.len()) };
        buf.truncate((n.max(0) as usize).min(buf.len() - 1));
        String::from_utf8_lossy(&buf).into_owned()
    }

    /// Device bytes held by this state (`fdtd_b200_memory_in_use`).
    pub fn memory_in_use(&self) -> u64 {
        unsafe { ffi::fdtd_b200_memory_in_use(self.h) }
    }

    /// Block until everything enqueued so far has run (`fdtd_b200_sync`).
    pub fn sync(&self) -> Result<(), B200Error> {
        check_handle(self.h, unsafe { ffi::fdtd_b200_sync(self.h) })
    }
}

impl Time for FdtdB200SolverState {
    fn time(&self) -> f64 {
        unsafe { ffi::fdtd_b200_time(self.h) }
    }

    fn tick(&self) -> usize {
        unsafe { ffi::fdtd_b200_tick(self.h) as usize }
    }
}

impl SolverInstance for FdtdB200SolverInstance {
    type State = FdtdB200SolverState;
    type UpdatePass<'a>
        = FdtdB200UpdatePass<'a>
    where
        Self: 'a;

    fn create_state(&self) -> FdtdB200SolverState {
        let mut h: ffi::fdtd_b200_handle = std::ptr::null_mut();
        let rc = unsafe {
            ffi::fdtd_b200_create(
                &self.cfg,
                &self.constants,
                self.palette.as_ptr(),
                self.palette.len() as u32,
                self.ids.as_ptr(),
                &mut h,
            )
        };
        // `create_state` is infallible in the trait and the runner `expect`s `create_instance` (runner.rs:352); a missing GPU
        // or an allocation failure surfaces here, loudly — there is no CPU fallback to fall back to.
        assert_eq!(rc, ffi::FDTD_B200_OK, "fdtd_b200_create failed: {}", global_error());
        let state = FdtdB200SolverState { h };
        for (axis, profile) in self.pml.iter().enumerate() {
            if let Some(profile) = profile {
                // one (b, c) set for the E and the H half-step = the reference's single PmlCoefficients per cell
                let rc = unsafe {
                    ffi::fdtd_b200_set_pml(
                        state.h,
                        axis as c_int,
                        profile.b.as_ptr(),
                        profile.c.as_ptr(),
                        std::ptr::null(),
                        std::ptr::null(),
                    )
                };
                assert_eq!(rc, ffi::FDTD_B200_OK, "fdtd_b200_set_pml failed: {}", last_error(state.h));
            }
        }
        tracing::debug!(path = %state.step_path(), bytes = state.memory_in_use(), "fdtd-b200 state created");
        state
    }

    fn begin_update<'a>(&'a self, state: &'a mut FdtdB200SolverState) -> FdtdB200UpdatePass<'a> {
        // forcing is per step: a new pass starts with none (cpu/mod.rs:370-384 clears `source_field`)
        FdtdB200UpdatePass {
            instance: self,
            state,
            sources: Vec::new(),
        }
    }
}

/// `UpdatePass` + `UpdatePassForcing` (lib.rs:81-90).
#[derive(Debug)]
pub struct FdtdB200UpdatePass<'a> {
    instance: &'a FdtdB200SolverInstance,
    state: &'a mut FdtdB200SolverState,
    sources: Vec<ffi::fdtd_b200_source>,
}

impl UpdatePassForcing<Point3<usize>> for FdtdB200UpdatePass<'_> {
    fn set_forcing(&mut self, point: &Point3<usize>, value: &SourceValues) {
        // the reference panics on a point outside the lattice (cpu/mod.rs:392-396, wgpu/mod.rs:342-346)
        let cell = self
            .instance
            .lattice
            .index(point)
            .unwrap_or_else(|| panic!("set_forcing called with invalid point: {point:?}"));
        // f64 → f32 once, as wgpu/mod.rs:626-635; later entries for the same cell win inside the library (cpu/mod.rs:399-408)
        self.sources.push(ffi::fdtd_b200_source {
            cell: cell as u64,
            j: [value.j.x as f32, value.j.y as f32, value.j.z as f32],
            m: [value.m.x as f32, value.m.y as f32, value.m.z as f32],
        });
    }
}

impl UpdatePass for FdtdB200UpdatePass<'_> {
    fn finish(self) {
        let h = self.state.h;
        // H half-step, E half-step on the new H, tick += 1, time += dt (cpu/mod.rs:416-515): enqueued, not awaited
        let rc = unsafe { ffi::fdtd_b200_set_sources(h, self.sources.len() as u32, self.sources.as_ptr()) };
        assert_eq!(rc, ffi::FDTD_B200_OK, "fdtd_b200_set_sources failed: {}", last_error(h));
        let rc = unsafe { ffi::fdtd_b200_step(h, 1) };
        assert_eq!(rc, ffi::FDTD_B200_OK, "fdtd_b200_step failed: {}", last_error(h));
    }
}
