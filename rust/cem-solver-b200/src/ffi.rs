//! 1:1 declarations of `include/fdtd_b200.h` (ABI version 2).  Checked against the header by
//! `tests/test_rust_shim.py`: every function's name, arity and argument types, every `#[repr(C)]` struct's field order and
//! size.  Nothing here is safe to call directly; `lib.rs` wraps it.
#![allow(non_camel_case_types, clippy::too_many_arguments)]

use std::os::raw::{c_char, c_int, c_void};

pub const FDTD_B200_ABI_VERSION: c_int = 2;

pub const FDTD_B200_OK: c_int = 0;
pub const FDTD_B200_EINVAL: c_int = 1;
pub const FDTD_B200_ENOMEM: c_int = 2;
pub const FDTD_B200_ECUDA: c_int = 3;
pub const FDTD_B200_ESTATE: c_int = 4;

pub const FDTD_B200_FIELD_E: c_int = 0;
pub const FDTD_B200_FIELD_H: c_int = 1;

pub const FDTD_B200_EXACT_ARITHMETIC: u32 = 1 << 0;
pub const FDTD_B200_FORCE_GENERIC_KERNELS: u32 = 1 << 1;
pub const FDTD_B200_FUSED_STEP: u32 = 1 << 2;
pub const FDTD_B200_IN_PLACE_STEP: u32 = 1 << 3;

pub const FDTD_B200_COLORMAP_CPU: c_int = 0;
pub const FDTD_B200_COLORMAP_LINEAR: c_int = 1;
pub const FDTD_B200_COLORMAP_APP_Z: c_int = 2;

pub const FDTD_B200_SHAPE_BALL: u32 = 0;
pub const FDTD_B200_SHAPE_CUBOID: u32 = 1;
pub const FDTD_B200_SHAPE_CYLINDER: u32 = 2;

pub const FDTD_B200_P2P_BLOB_BYTES: usize = 2048;
pub const FDTD_B200_STATIC_SCRATCH_BYTES: usize = 8 << 20;

/// Opaque `struct fdtd_b200_solver`; `fdtd_b200_handle` is a pointer to it.
#[repr(C)]
pub struct fdtd_b200_solver {
    _private: [u8; 0],
}
pub type fdtd_b200_handle = *mut fdtd_b200_solver;

/// cem-solver/src/material.rs:105-116 (same field order, same spelling).
#[repr(C)]
#[derive(Clone, Copy, Debug, Default, PartialEq)]
pub struct fdtd_b200_material {
    pub relative_permeability: f64,
    pub magnetic_conductivity: f64,
    pub relative_permittivity: f64,
    pub eletrical_conductivity: f64,
}

/// cem-solver/src/material.rs:21-25
#[repr(C)]
#[derive(Clone, Copy, Debug, Default, PartialEq)]
pub struct fdtd_b200_physical_constants {
    pub vacuum_permittivity: f64,
    pub vacuum_permeability: f64,
}

/// cem-solver/src/fdtd/util.rs:91-96
#[repr(C)]
#[derive(Clone, Copy, Debug, Default, PartialEq)]
pub struct fdtd_b200_update_coefficients {
    pub c_a: f64,
    pub c_b: f64,
    pub d_a: f64,
    pub d_b: f64,
}

/// cem-solver/src/fdtd/wgpu/mod.rs:661-679
#[repr(C)]
#[derive(Clone, Copy, Debug, Default, PartialEq)]
pub struct fdtd_b200_coefficients_f32 {
    pub c_a: f32,
    pub c_b: f32,
    pub d_a: f32,
    pub d_b: f32,
}

/// Replaces `FdtdSolverConfig` + `Resolution` (cem-solver/src/fdtd/mod.rs:18-22, 64-67).
#[repr(C)]
#[derive(Clone, Copy, Debug, Default, PartialEq)]
pub struct fdtd_b200_config {
    pub nx: u64,
    pub ny: u64,
    pub nz: u64,
    pub dx: [f64; 3],
    pub dt: f64,
    pub device: i32,
    pub flags: u32,
    pub z_begin: u64,
    pub z_end: u64,
    pub probe_ring_steps: u32,
    pub reserved: u32,
}

/// ≙ `SourceData` (cem-solver/src/fdtd/wgpu/mod.rs:617-624) with a 64-bit cell index.
#[repr(C)]
#[derive(Clone, Copy, Debug, Default, PartialEq)]
pub struct fdtd_b200_source {
    pub cell: u64,
    pub j: [f32; 3],
    pub m: [f32; 3],
}

/// cem-solver/src/fdtd/pml.rs:13-20 without the normal.
#[repr(C)]
#[derive(Clone, Copy, Debug, Default, PartialEq)]
pub struct fdtd_b200_graded_pml {
    pub m: f64,
    pub m_a: f64,
    pub sigma_max: f64,
    pub kappa_max: f64,
    pub a_max: f64,
}

/// One collider of the application's scene (cem-app/src/solver/runner.rs:606-631).
#[repr(C)]
#[derive(Clone, Copy, Debug, Default, PartialEq)]
pub struct fdtd_b200_shape {
    pub kind: u32,
    pub label: u32,
    pub params: [f32; 3],
    pub rot: [f32; 9],
    pub translation: [f32; 3],
    pub reserved: f32,
}

#[link(name = "fdtd_b200")]
unsafe extern "C" {
    // ---- library
    pub fn fdtd_b200_abi_version() -> c_int;
    pub fn fdtd_b200_device_count() -> c_int;
    pub fn fdtd_b200_global_error() -> *const c_char;

    // ---- host-side model helpers
    pub fn fdtd_b200_update_coefficients_new(
        dt: f64,
        constants: *const fdtd_b200_physical_constants,
        material: *const fdtd_b200_material,
        out: *mut fdtd_b200_update_coefficients,
    ) -> c_int;
    pub fn fdtd_b200_index_materials(
        materials: *const fdtd_b200_material,
        n_cells: u64,
        ids_out: *mut u32,
        palette_out: *mut fdtd_b200_material,
        palette_cap: u32,
        n_palette_out: *mut u32,
    ) -> c_int;
    pub fn fdtd_b200_canonicalize_material_index(
        ids_in: *const u32,
        n_cells: u64,
        palette_in: *const fdtd_b200_material,
        n_palette_in: u32,
        ids_out: *mut u32,
        palette_out: *mut fdtd_b200_material,
        palette_cap: u32,
        n_palette_out: *mut u32,
    ) -> c_int;
    pub fn fdtd_b200_memory_required(cfg: *const fdtd_b200_config, index_bytes: u32) -> usize;
    pub fn fdtd_b200_choose_step_mode(
        cfg: *const fdtd_b200_config,
        index_bytes: u32,
        memory_limit: u64,
        flags_out: *mut u32,
        required_out: *mut u64,
    ) -> c_int;

    // ---- instance / state
    pub fn fdtd_b200_create(
        cfg: *const fdtd_b200_config,
        constants: *const fdtd_b200_physical_constants,
        palette: *const fdtd_b200_material,
        n_palette: u32,
        material_ids: *const u32,
        out: *mut fdtd_b200_handle,
    ) -> c_int;
    pub fn fdtd_b200_create_from_table(
        cfg: *const fdtd_b200_config,
        table: *const fdtd_b200_coefficients_f32,
        n_table: u32,
        material_ids: *const u32,
        out: *mut fdtd_b200_handle,
    ) -> c_int;
    pub fn fdtd_b200_destroy(h: fdtd_b200_handle);
    pub fn fdtd_b200_last_error(h: fdtd_b200_handle) -> *const c_char;
    pub fn fdtd_b200_reset(h: fdtd_b200_handle) -> c_int;
    pub fn fdtd_b200_set_stream(h: fdtd_b200_handle, cuda_stream: *mut c_void) -> c_int;
    pub fn fdtd_b200_reset_stream(h: fdtd_b200_handle) -> c_int;
    pub fn fdtd_b200_sync(h: fdtd_b200_handle) -> c_int;
    pub fn fdtd_b200_tick(h: fdtd_b200_handle) -> u64;
    pub fn fdtd_b200_time(h: fdtd_b200_handle) -> f64;
    pub fn fdtd_b200_set_time(h: fdtd_b200_handle, tick: u64, time: f64) -> c_int;

    // ---- CPML
    pub fn fdtd_b200_pml_coefficients(
        dt: f64,
        constants: *const fdtd_b200_physical_constants,
        sigma: f64,
        kappa: f64,
        a: f64,
        normal: *const f64,
        b_out: *mut f64,
        c_out: *mut f64,
    ) -> c_int;
    pub fn fdtd_b200_pml_coefficients_graded(
        dt: f64,
        constants: *const fdtd_b200_physical_constants,
        g: *const fdtd_b200_graded_pml,
        depth: f64,
        normal: *const f64,
        b_out: *mut f64,
        c_out: *mut f64,
    ) -> c_int;
    pub fn fdtd_b200_set_pml(
        h: fdtd_b200_handle,
        axis: c_int,
        b_e: *const f64,
        c_e: *const f64,
        b_h: *const f64,
        c_h: *const f64,
    ) -> c_int;
    pub fn fdtd_b200_has_pml(h: fdtd_b200_handle) -> c_int;
    pub fn fdtd_b200_pml_in_kernel(h: fdtd_b200_handle) -> c_int;
    pub fn fdtd_b200_pml_state_floats(h: fdtd_b200_handle) -> u64;
    pub fn fdtd_b200_read_pml_state(h: fdtd_b200_handle, out: *mut f32) -> c_int;
    pub fn fdtd_b200_write_pml_state(h: fdtd_b200_handle, r#in: *const f32) -> c_int;

    // ---- introspection
    pub fn fdtd_b200_memory_in_use(h: fdtd_b200_handle) -> u64;
    pub fn fdtd_b200_step_path(h: fdtd_b200_handle, buf: *mut c_char, cap: usize) -> c_int;
    pub fn fdtd_b200_index_bytes(h: fdtd_b200_handle) -> c_int;
    pub fn fdtd_b200_uses_vec4(h: fdtd_b200_handle) -> c_int;
    pub fn fdtd_b200_uses_bulk(h: fdtd_b200_handle) -> c_int;
    pub fn fdtd_b200_uses_fused(h: fdtd_b200_handle) -> c_int;
    pub fn fdtd_b200_kernel_launches(h: fdtd_b200_handle) -> u64;
    pub fn fdtd_b200_set_profiling(h: fdtd_b200_handle, enabled: c_int) -> c_int;
    pub fn fdtd_b200_get_profile(
        h: fdtd_b200_handle,
        which: c_int,
        total_ms: *mut f64,
        n_launches: *mut u64,
        n_cells: *mut u64,
    ) -> c_int;
    pub fn fdtd_b200_timer_start(h: fdtd_b200_handle) -> c_int;
    pub fn fdtd_b200_timer_stop(h: fdtd_b200_handle, elapsed_ms: *mut f64) -> c_int;
    pub fn fdtd_b200_read_material_ids(h: fdtd_b200_handle, ids_out: *mut u32) -> c_int;
    pub fn fdtd_b200_read_coefficient_table(
        h: fdtd_b200_handle,
        out: *mut fdtd_b200_coefficients_f32,
        cap: u32,
        n_out: *mut u32,
    ) -> c_int;

    // ---- update pass
    pub fn fdtd_b200_set_sources(h: fdtd_b200_handle, n: u32, sources: *const fdtd_b200_source) -> c_int;
    pub fn fdtd_b200_set_source_table(
        h: fdtd_b200_handle,
        n_src: u32,
        cells: *const u64,
        n_steps: u64,
        values: *const f32,
    ) -> c_int;
    pub fn fdtd_b200_step(h: fdtd_b200_handle, n_steps: u64) -> c_int;
    pub fn fdtd_b200_update_h(h: fdtd_b200_handle, z_lo: u64, z_hi: u64) -> c_int;
    pub fn fdtd_b200_update_e(h: fdtd_b200_handle, z_lo: u64, z_hi: u64) -> c_int;
    pub fn fdtd_b200_end_step(h: fdtd_b200_handle) -> c_int;
    pub fn fdtd_b200_fused_update(h: fdtd_b200_handle, z_lo: u64, z_hi: u64) -> c_int;
    pub fn fdtd_b200_fused_commit(h: fdtd_b200_handle) -> c_int;
    pub fn fdtd_b200_halo_plane(
        h: fdtd_b200_handle,
        which: c_int,
        device_ptr: *mut *mut c_void,
        n_floats: *mut u64,
    ) -> c_int;
    pub fn fdtd_b200_halo_copy(
        dst: fdtd_b200_handle,
        which_dst: c_int,
        src: fdtd_b200_handle,
        which_src: c_int,
    ) -> c_int;

    // ---- NVLink peer-memory halo transport
    pub fn fdtd_b200_p2p_export(h: fdtd_b200_handle, blob: *mut c_void) -> c_int;
    pub fn fdtd_b200_p2p_connect(
        h: fdtd_b200_handle,
        lower_blob: *const c_void,
        upper_blob: *const c_void,
    ) -> c_int;
    pub fn fdtd_b200_p2p_refresh(h: fdtd_b200_handle) -> c_int;
    pub fn fdtd_b200_p2p_step(h: fdtd_b200_handle, n_steps: u64) -> c_int;
    pub fn fdtd_b200_p2p_connected(h: fdtd_b200_handle) -> c_int;

    // ---- scene -> material lattice
    pub fn fdtd_b200_rasterize_scene(
        device: c_int,
        dims: *const u64,
        z_begin: u64,
        z_end: u64,
        solver_to_world: *const f64,
        shapes: *const fdtd_b200_shape,
        n_shapes: u32,
        default_label: u32,
        ids_out: *mut u32,
    ) -> c_int;
    pub fn fdtd_b200_set_materials_from_scene(
        h: fdtd_b200_handle,
        solver_to_world: *const f64,
        shapes: *const fdtd_b200_shape,
        n_shapes: u32,
        default_label: u32,
    ) -> c_int;

    // ---- field access
    pub fn fdtd_b200_read_field(
        h: fdtd_b200_handle,
        component: c_int,
        lo: *const u64,
        hi: *const u64,
        out: *mut f32,
    ) -> c_int;
    pub fn fdtd_b200_read_field_static(
        h: fdtd_b200_handle,
        component: c_int,
        lo: *const u64,
        hi: *const u64,
        out: *mut f32,
    ) -> c_int;
    pub fn fdtd_b200_write_field(
        h: fdtd_b200_handle,
        component: c_int,
        lo: *const u64,
        hi: *const u64,
        r#in: *const f32,
    ) -> c_int;
    pub fn fdtd_b200_device_field(h: fdtd_b200_handle, c: c_int, device_ptr: *mut *mut c_void) -> c_int;
    pub fn fdtd_b200_row_pitch(h: fdtd_b200_handle) -> u64;
    pub fn fdtd_b200_register_probes(h: fdtd_b200_handle, n: u32, cells: *const u64) -> c_int;
    pub fn fdtd_b200_read_probes(h: fdtd_b200_handle, first_tick: u64, n_steps: u64, out: *mut f32) -> c_int;

    // ---- projection
    pub fn fdtd_b200_project(
        h: fdtd_b200_handle,
        component: c_int,
        projection: *const f32,
        color_map: *const f32,
        colormap_mode: c_int,
        w: u32,
        hgt: u32,
        out_rgba: *mut u8,
        rgba_is_device: c_int,
    ) -> c_int;
}
