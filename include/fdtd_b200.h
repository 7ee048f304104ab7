/*
 * fdtd_b200 — C ABI of the B200-native (sm_100a) FDTD time-stepping engine that replaces
 * cem-solver's `fdtd::cpu` / `fdtd::wgpu` backends behind the reference's solver traits
 * (cem-solver/src/lib.rs:28-142, cem-solver/src/project.rs:75-151).
 *
 * Everything crossing this boundary is plain C: integers, doubles, host pointers and opaque
 * handles.  No call aborts or throws; every call returns FDTD_B200_OK (0) or an error code and
 * leaves a message retrievable through fdtd_b200_last_error().  There is no CPU fallback: when
 * no CUDA device is usable fdtd_b200_create() fails with FDTD_B200_ECUDA.
 *
 * One handle = reference `Instance` + `State` merged (cpu/mod.rs:225-232, 299-307): it owns the
 * material index lattice, the coefficient table, the six SoA fp32 field components, sources,
 * probes, tick and time.  A handle is single-caller (the reference drives it from one "solver"
 * thread, cem-app/src/solver/runner.rs:471) but may be created on one thread and used on another.
 *
 * Linear cell index = x + nx*(y + ny*z)   (cem-solver/src/fdtd/strider.rs:47-53, 147-154).
 */
#ifndef FDTD_B200_H
#define FDTD_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define FDTD_B200_ABI_VERSION 2

typedef struct fdtd_b200_solver* fdtd_b200_handle;

enum {
    FDTD_B200_OK = 0,
    FDTD_B200_EINVAL = 1, /* bad argument; an out-of-lattice source point maps to the reference's panic
                             (cpu/mod.rs:392-396, wgpu/mod.rs:342-346) in the Rust shim */
    FDTD_B200_ENOMEM = 2,
    FDTD_B200_ECUDA = 3, /* CUDA runtime error or no usable device */
    FDTD_B200_ESTATE = 4 /* call not valid in the handle's current state */
};

/* cem-solver/src/lib.rs:139-142 `FieldComponent` */
enum { FDTD_B200_FIELD_E = 0, FDTD_B200_FIELD_H = 1 };

/* fdtd_b200_config.flags */
enum {
    /* Arithmetic of update.wgsl:89,127 reproduced operation by operation: true division by dx, no
       FMA contraction, `+ psi(=0)` kept.  Bit-identical to the fp32 oracle.  Without the flag the
       kernels multiply by 1/dx and let the compiler contract to FMA (documented in DESIGN.md). */
    FDTD_B200_EXACT_ARITHMETIC = 1u << 0,
    /* Force the one-thread-per-cell kernels even when nx % 4 == 0 (debugging / bisecting). */
    FDTD_B200_FORCE_GENERIC_KERNELS = 1u << 1,
    /* Accepted for source compatibility with ABI 1 and ignored: the fused single-pass step is the DEFAULT since ABI 2. */
    FDTD_B200_FUSED_STEP = 1u << 2,
    /* Step kernel selection.  `SolverBackend::create_instance` (lib.rs:28-45) has no knob for it, so the library decides:
       by default fdtd_b200_step() runs ONE fused kernel per step (H and E half-steps in a single pass over the lattice,
       49 B instead of 74 B of HBM traffic per cell update) on ping-pong field buffers, i.e. +24 B/cell of device memory,
       whenever the lattice is eligible (row pitch >= 128 cells).  This flag opts out: two half-step kernels updating
       the fields in place (24 B/cell of field memory).  Results are bit-identical either way.  A caller with a memory
       budget (cem-app/src/solver/runner.rs:304-330 `memory_limit`) asks fdtd_b200_choose_step_mode() which of the two
       fits; fdtd_b200_step_path() reports what a handle runs and why. */
    FDTD_B200_IN_PLACE_STEP = 1u << 3
};

/* cem-solver/src/material.rs:105-116 (same field order). */
typedef struct fdtd_b200_material {
    double relative_permeability;  /* mu_r    */
    double magnetic_conductivity;  /* sigma_m */
    double relative_permittivity;  /* eps_r   */
    double eletrical_conductivity; /* sigma (spelling as in the reference) */
} fdtd_b200_material;

/* cem-solver/src/material.rs:21-25 */
typedef struct fdtd_b200_physical_constants {
    double vacuum_permittivity;
    double vacuum_permeability;
} fdtd_b200_physical_constants;

/* cem-solver/src/fdtd/util.rs:91-96 — derived in f64 ... */
typedef struct fdtd_b200_update_coefficients {
    double c_a, c_b, d_a, d_b;
} fdtd_b200_update_coefficients;

/* ... and cast once to fp32 for the device (wgpu/mod.rs:661-679). */
typedef struct fdtd_b200_coefficients_f32 {
    float c_a, c_b, d_a, d_b;
} fdtd_b200_coefficients_f32;

/* Replaces FdtdSolverConfig + Resolution (fdtd/mod.rs:18-22, 64-67).  Lattice dims are integers
 * because `ceil(size/dx)` of inexact ratios is the caller's business (fdtd/mod.rs:25-29). */
typedef struct fdtd_b200_config {
    uint64_t nx, ny, nz; /* global lattice */
    double dx[3];        /* Resolution::spatial */
    double dt;           /* Resolution::temporal */
    int32_t device;      /* CUDA device ordinal */
    uint32_t flags;
    /* z-slab owned by this handle: global planes [z_begin, z_end).  0,0 means the whole lattice.
       A slab with z_begin > 0 / z_end < nz keeps one ghost plane below / above that the caller
       (or fdtd_b200_exchange_*) fills from the neighbouring rank. */
    uint64_t z_begin, z_end;
    uint32_t probe_ring_steps; /* capacity of the probe ring in steps; 0 = default (4096) */
    uint32_t reserved;
} fdtd_b200_config;

/* Per-step forcing of one cell ≙ `SourceData` (wgpu/mod.rs:617-624) with a 64-bit cell index
 * (the reference's u32 index cannot address 2048^3).  j feeds the E half-step, m the H half-step
 * of the same step (cpu/mod.rs:445-451, 487-493). */
typedef struct fdtd_b200_source {
    uint64_t cell; /* global linear index */
    float j[3];
    float m[3];
} fdtd_b200_source;

/* ---- library ------------------------------------------------------------------------------ */

int fdtd_b200_abi_version(void);
/* Number of CUDA devices visible (0 if none / no driver). */
int fdtd_b200_device_count(void);
/* Thread-local message of the last failing call made without a handle (create, helpers). */
const char* fdtd_b200_global_error(void);

/* ---- host-side model helpers (f64, no device needed) --------------------------------------- */

/* `UpdateCoefficients::new` (fdtd/util.rs:99-128).  EINVAL where the reference's NaN asserts fire. */
int fdtd_b200_update_coefficients_new(double dt, const fdtd_b200_physical_constants* constants,
                                      const fdtd_b200_material* material, fdtd_b200_update_coefficients* out);

/* The per-cell coefficient buffer compressed into index + table (README.md:25-28 TODO).
 * `materials` holds one record per cell (n_cells) in linear-index order — what the reference's
 * per-cell `DomainDescription::material(point)` loop yields (cpu/mod.rs:244-264).  Ids are handed
 * out in first-occurrence order, key = bit pattern of the four doubles.  ids_out: n_cells u32.
 * palette_out: up to palette_cap records.  *n_palette_out: number of distinct materials (may
 * exceed palette_cap; then only the first palette_cap are written and EINVAL is returned). */
int fdtd_b200_index_materials(const fdtd_b200_material* materials, uint64_t n_cells, uint32_t* ids_out,
                              fdtd_b200_material* palette_out, uint32_t palette_cap, uint32_t* n_palette_out);

/* Same canonical numbering starting from an already-indexed lattice (ids_in may be any labelling
 * of `palette_in`; equal materials under different labels are merged).  ids_in/ids_out may alias. */
int fdtd_b200_canonicalize_material_index(const uint32_t* ids_in, uint64_t n_cells, const fdtd_b200_material* palette_in,
                                          uint32_t n_palette_in, uint32_t* ids_out, fdtd_b200_material* palette_out,
                                          uint32_t palette_cap, uint32_t* n_palette_out);

/* `SolverBackend::memory_required` (lib.rs:41-44; cpu/mod.rs:217-221; wgpu/mod.rs:129-133): device bytes of the fields and
 * the material index a handle for `cfg` will allocate with an index of `index_bytes` (1 or 2) per cell, in the step mode
 * cfg->flags selects (fused: two field sets unless FDTD_B200_IN_PLACE_STEP is set or the lattice is not eligible).  This is
 * the per-cell part; probe rings, PML psi arrays and forcing tables come on top (fdtd_b200_memory_in_use reports the total
 * of a live handle). */
size_t fdtd_b200_memory_required(const fdtd_b200_config* cfg, uint32_t index_bytes);
/* What cem-app/src/solver/runner.rs:304-330 does with `memory_limit`, for a backend with two step modes: the fused step if
 * its memory fits `memory_limit` bytes (0 = no limit), else the in-place two-kernel step, else ENOMEM.  *flags_out =
 * cfg->flags with FDTD_B200_IN_PLACE_STEP set or cleared accordingly, *required_out = fdtd_b200_memory_required of that
 * mode (also on ENOMEM: the smaller of the two). */
int fdtd_b200_choose_step_mode(const fdtd_b200_config* cfg, uint32_t index_bytes, uint64_t memory_limit, uint32_t* flags_out,
                               uint64_t* required_out);

/* ---- instance / state ---------------------------------------------------------------------- */

/* `SolverBackend::create_instance` + `SolverInstance::create_state` (lib.rs:28-45, 70-79).
 * material_ids: host, one u32 per cell of THIS handle's slab ([z_begin,z_end) planes, x fastest);
 * NULL = every cell uses palette[0].  The device copy is u8 when n_palette <= 256, else u16
 * (n_palette <= 65536).  Coefficients are derived in f64 and cast once to fp32. */
int fdtd_b200_create(const fdtd_b200_config* cfg, const fdtd_b200_physical_constants* constants,
                     const fdtd_b200_material* palette, uint32_t n_palette, const uint32_t* material_ids,
                     fdtd_b200_handle* out);
/* Same, from a caller-built fp32 coefficient table. */
int fdtd_b200_create_from_table(const fdtd_b200_config* cfg, const fdtd_b200_coefficients_f32* table, uint32_t n_table,
                                const uint32_t* material_ids, fdtd_b200_handle* out);
void fdtd_b200_destroy(fdtd_b200_handle h);
const char* fdtd_b200_last_error(fdtd_b200_handle h);

/* A fresh `create_state` (cpu/mod.rs:310-322): zero fields, clear sources, tick = 0, time = 0. */
int fdtd_b200_reset(fdtd_b200_handle h);
/* Run subsequent work on this CUDA stream (a cudaStream_t; NULL means the legacy default stream, which is
 * what torch's default stream is).  reset_stream goes back to the handle's own non-blocking stream. */
int fdtd_b200_set_stream(fdtd_b200_handle h, void* cuda_stream);
int fdtd_b200_reset_stream(fdtd_b200_handle h);
int fdtd_b200_sync(fdtd_b200_handle h);

/* `Time` (lib.rs:92-95).  time is accumulated `time += dt` in f64 (cpu/mod.rs:513). */
uint64_t fdtd_b200_tick(fdtd_b200_handle h);
double fdtd_b200_time(fdtd_b200_handle h);
/* Restore tick/time of a checkpointed state (fields come back through fdtd_b200_write_field).  Drops pending forcing
 * rows and restarts the probe ring at `tick`. */
int fdtd_b200_set_time(fdtd_b200_handle h, uint64_t tick, double time);

/* ---- CPML ------------------------------------------------------------------------
 * The reference computes PML coefficients (fdtd/pml.rs:34-86), keeps a per-cell indirection + ψ storage
 * (cpu/mod.rs:242-255, 649-679) and never runs the recursion (`psi = zeros`, cpu/mod.rs:419-425, 454, 498): parity unpinned.
 * Here the recursion runs, for separable PMLs: per axis a profile (b[i], c[i]) over the GLOBAL coordinate along that axis,
 * one for the E half-step and one for the H half-step (half a cell apart on the Yee grid; the reference has one
 * `PmlCoefficients` per cell for both — pass NULL for b_h/c_h to get that).  Cells whose c is non-zero on an axis carry two
 * ψ values per half-step for that axis; psi = ±curl(G_x, G_y, G_z) enters the update where the reference has its `+ psi`. */

/* cem-solver/src/fdtd/pml.rs:13-20 without the normal. */
typedef struct fdtd_b200_graded_pml {
    double m, m_a, sigma_max, kappa_max, a_max;
} fdtd_b200_graded_pml;

/* `PmlCoefficients::new` (pml.rs:34-56): scalar b, c times `normal` (host, f64). */
int fdtd_b200_pml_coefficients(double dt, const fdtd_b200_physical_constants* constants, double sigma, double kappa, double a,
                               const double normal[3], double b_out[3], double c_out[3]);
/* `PmlCoefficients::new_graded` (pml.rs:58-86); depth in [0, 1]. */
int fdtd_b200_pml_coefficients_graded(double dt, const fdtd_b200_physical_constants* constants, const fdtd_b200_graded_pml* g, double depth,
                                      const double normal[3], double b_out[3], double c_out[3]);
/* Profiles of one axis (arrays of the global extent of that axis); b_e == NULL clears the axis.  ψ starts at zero.
 * Fused-step handles run the recursion inside the fused kernel when the x slabs fit its tiling (at most 32 PML
 * x-coordinates per 128-cell tile; FDTD_B200_PML_PASSES=1 in the environment forces the
 * alternative), else — whole lattices only — as extra passes after it.  Fused z-slab handles need the in-kernel form (else
 * ESTATE: use two-kernel handles) and the peer-memory transport: a slab recomputes H' of the plane above it and receives
 * that plane's psi_H from the slab above with the boundary planes (fdtd_b200_p2p_*; set the PML on every slab BEFORE
 * fdtd_b200_p2p_export, the blobs carry its layout). */
int fdtd_b200_set_pml(fdtd_b200_handle h, int axis, const double* b_e, const double* c_e, const double* b_h, const double* c_h);
int fdtd_b200_has_pml(fdtd_b200_handle h);
int fdtd_b200_pml_in_kernel(fdtd_b200_handle h);   /* 1 when the fused kernel itself advances ψ (no PML passes) */
/* ψ state for checkpoints: all ψ arrays of this handle, concatenated. */
uint64_t fdtd_b200_pml_state_floats(fdtd_b200_handle h);
int fdtd_b200_read_pml_state(fdtd_b200_handle h, float* out);
int fdtd_b200_write_pml_state(fdtd_b200_handle h, const float* in);

/* Introspection used by tests and the bench. */
/* Device bytes this handle holds right now (fields, index, table, forcing, probe ring, psi, staging) — cudaMalloc sizes. */
uint64_t fdtd_b200_memory_in_use(fdtd_b200_handle h);
/* Which kernels fdtd_b200_step() launches on this handle and why, e.g. "fused: k_step_fused<fast,u8>" or
 * "two-kernel (bulk): in-place step requested (FDTD_B200_IN_PLACE_STEP)" or "two-kernel (vec4): row pitch 64 < 128 cells".
 * Writes at most cap bytes including the terminating NUL; returns the full length. */
int fdtd_b200_step_path(fdtd_b200_handle h, char* buf, size_t cap);
int fdtd_b200_index_bytes(fdtd_b200_handle h);   /* 1 or 2 */
int fdtd_b200_uses_vec4(fdtd_b200_handle h);     /* 1 when the float4 z-streaming kernels run */
int fdtd_b200_uses_bulk(fdtd_b200_handle h);     /* 1 when the TMA bulk-copy staged kernels run */
int fdtd_b200_uses_fused(fdtd_b200_handle h);    /* 1 when fdtd_b200_step runs the fused single-pass kernel */
uint64_t fdtd_b200_kernel_launches(fdtd_b200_handle h); /* kernels launched so far by this handle */
/* Measurement: with profiling on, every H / E half-step launch is bracketed by CUDA events on the
 * handle's stream.  get_profile(which = 0 H | 1 E) blocks and returns the summed device time, the
 * number of launches and the cells they updated since profiling was switched on. */
int fdtd_b200_set_profiling(fdtd_b200_handle h, int enabled);
int fdtd_b200_get_profile(fdtd_b200_handle h, int which, double* total_ms, uint64_t* n_launches, uint64_t* n_cells);
/* CUDA-event stopwatch on the handle's stream: start synchronises then records; stop records,
 * waits and returns the device time between the two. */
int fdtd_b200_timer_start(fdtd_b200_handle h);
int fdtd_b200_timer_stop(fdtd_b200_handle h, double* elapsed_ms);
/* Device copy of the material index back to the host as u32 (slab order). */
int fdtd_b200_read_material_ids(fdtd_b200_handle h, uint32_t* ids_out);
int fdtd_b200_read_coefficient_table(fdtd_b200_handle h, fdtd_b200_coefficients_f32* out, uint32_t cap, uint32_t* n_out);

/* ---- update pass --------------------------------------------------------------------------- */

/* `begin_update` + `set_forcing`* (cpu/mod.rs:370-409): the forcing of the NEXT step only.  Later
 * entries for the same cell win (CPU backend semantics).  Cells are global linear indices; entries
 * outside the lattice → EINVAL and nothing is registered; entries outside this handle's slab are
 * ignored.  Cleared after the step, like the reference clears `source_field`. */
int fdtd_b200_set_sources(fdtd_b200_handle h, uint32_t n, const fdtd_b200_source* sources);

/* Batched forcing for fdtd_b200_step(h, n>1): values[s][k][0..2] = j, [3..5] = m of source k at the
 * s-th step from now; row s is consumed by the s-th following step; steps beyond n_steps run
 * unforced.  Duplicate cells: the later column wins.  n_src = 0 clears the table. */
int fdtd_b200_set_source_table(fdtd_b200_handle h, uint32_t n_src, const uint64_t* cells, uint64_t n_steps,
                               const float* values);

/* `UpdatePass::finish` × n (cpu/mod.rs:416-515): H half-step, E half-step on the new H, tick += 1,
 * time += dt.  Enqueues and returns; errors surface at the next blocking call.  On a slab handle
 * this does NOT exchange halos — use the phase calls below or fdtd_b200_step with a single slab. */
int fdtd_b200_step(fdtd_b200_handle h, uint64_t n_steps);

/* Phases of one step for slab-decomposed runs: update local planes [z_lo, z_hi) of H or E. */
int fdtd_b200_update_h(fdtd_b200_handle h, uint64_t z_lo, uint64_t z_hi);
int fdtd_b200_update_e(fdtd_b200_handle h, uint64_t z_lo, uint64_t z_hi);
/* tick += 1, time += dt, consume one forcing row, advance the probe ring. */
int fdtd_b200_end_step(fdtd_b200_handle h);
/* Fused-step handles (fdtd_b200_uses_fused): update planes [z_lo, z_hi) with the single-pass kernel, reading the current
 * buffer set (incl. ghost planes) and writing the pending one; launches of disjoint ranges are independent.  commit
 * makes the pending set current (call once per step after every plane was updated, then fdtd_b200_end_step). */
int fdtd_b200_fused_update(fdtd_b200_handle h, uint64_t z_lo, uint64_t z_hi);
int fdtd_b200_fused_commit(fdtd_b200_handle h);

/* Halo planes (device pointers, nx*ny floats each).  which: */
enum {
    FDTD_B200_HALO_SEND_EX_TOP = 0, /* last owned plane of Ex — goes up after the E half-step */
    FDTD_B200_HALO_SEND_EY_TOP = 1,
    FDTD_B200_HALO_RECV_EX_LOW = 2, /* ghost plane below plane 0 */
    FDTD_B200_HALO_RECV_EY_LOW = 3,
    FDTD_B200_HALO_SEND_HX_BOTTOM = 4, /* first owned plane of Hx — goes down after the H half-step */
    FDTD_B200_HALO_SEND_HY_BOTTOM = 5,
    FDTD_B200_HALO_RECV_HX_HIGH = 6, /* ghost plane above the last plane */
    FDTD_B200_HALO_RECV_HY_HIGH = 7,
    /* fused-step slabs additionally ship their whole first plane down (the lower rank recomputes H' of it): */
    FDTD_B200_HALO_SEND_EX_BOTTOM = 8,
    FDTD_B200_HALO_SEND_EY_BOTTOM = 9,
    FDTD_B200_HALO_SEND_EZ_BOTTOM = 10,
    FDTD_B200_HALO_RECV_EX_HIGH = 11,
    FDTD_B200_HALO_RECV_EY_HIGH = 12,
    FDTD_B200_HALO_RECV_EZ_HIGH = 13,
    /* material index of the first owned plane / of the plane above the slab (exchanged once; n_floats = bytes/4) */
    FDTD_B200_HALO_SEND_MAT_BOTTOM = 14,
    FDTD_B200_HALO_RECV_MAT_HIGH = 15,
    /* OR-ed into a selector: the plane of the PENDING buffer set of a fused-step handle (what fdtd_b200_fused_update
       is writing and fdtd_b200_fused_commit will make current) */
    FDTD_B200_HALO_PENDING = 0x100
};
int fdtd_b200_halo_plane(fdtd_b200_handle h, int which, void** device_ptr, uint64_t* n_floats);
/* Single-process multi-GPU transport: copy halo plane `which_src` of `src` into `which_dst` of `dst`
 * (cudaMemcpyPeerAsync over NVLink when the handles live on different devices), ordered after the
 * work already enqueued on src's stream and before later work on dst's stream. */
int fdtd_b200_halo_copy(fdtd_b200_handle dst, int which_dst, fdtd_b200_handle src, int which_src);

/* ---- NVLink peer-memory halo transport (one process per GPU, no NCCL, no host synchronisation per step) ----
 * Fused-step slab handles only.  Every rank exports a blob describing its device allocations (CUDA IPC memory handles; raw
 * pointers when both handles live in one process), the caller moves the blobs between z-neighbours by any out-of-band
 * channel (torch.distributed all_gather in fdtd_b200/distributed.py), and connect maps the neighbours' ghost planes.
 * fdtd_b200_p2p_step then runs n leapfrog steps: the slab's boundary planes are updated first on a high-priority stream,
 * stored straight into the neighbours' ghost planes over NVLink (SURVEY §8e "primary" transport), followed by a release
 * flag in the neighbour's memory; the neighbour's next boundary launch acquires that flag on the device; the interior
 * planes run concurrently on the main stream.  All ranks must make the same sequence of p2p_refresh / p2p_step calls. */
#define FDTD_B200_P2P_BLOB_BYTES 2048
int fdtd_b200_p2p_export(fdtd_b200_handle h, void* blob /* FDTD_B200_P2P_BLOB_BYTES */);
/* lower_blob / upper_blob: blobs of the ranks owning the slab below / above (NULL where the lattice ends). */
int fdtd_b200_p2p_connect(fdtd_b200_handle h, const void* lower_blob, const void* upper_blob);
/* Push the CURRENT boundary planes (and, once, the material index of the first plane; with a PML also psi_H of the first
 * plane, into the lower neighbour's receive slot of the current tick's parity) into the neighbours' ghost planes — after
 * connect, reset, fdtd_b200_set_time, fdtd_b200_write_pml_state and after fields were written from outside.  Asynchronous:
 * ordered against the neighbours' steps by the same device-side flags, so no host barrier is needed as long as every rank
 * calls it at the same point.  (fdtd_b200_p2p_export allocates the psi receive buffer: export again, and re-connect the
 * neighbours, after changing the PML.) */
int fdtd_b200_p2p_refresh(fdtd_b200_handle h);
int fdtd_b200_p2p_step(fdtd_b200_handle h, uint64_t n_steps);
/* 1 when connect succeeded and p2p_step may be called. */
int fdtd_b200_p2p_connected(fdtd_b200_handle h);

/* ---- scene → material lattice --------------------------------------------------------------- */

/* One collider of the application's scene with the material label its entity carries.  Replaces the per-cell
 * `DomainDescription::material(point)` callback (cpu/mod.rs:244-264) as the application implements it: lattice point →
 * world (f64 matrix, cast to f32) → first collider containing the point, else the default material
 * (cem-app/src/solver/runner.rs:606-631, 909-918).  Primitives are parry3d's Ball / Cuboid / Cylinder (axis = local Y). */
enum { FDTD_B200_SHAPE_BALL = 0, FDTD_B200_SHAPE_CUBOID = 1, FDTD_B200_SHAPE_CYLINDER = 2 };
typedef struct fdtd_b200_shape {
    uint32_t kind;
    uint32_t label;       /* palette index written to the cells whose point lies inside */
    float params[3];      /* ball: radius | cuboid: half extents x,y,z | cylinder: half_height, radius */
    float rot[9];         /* world → local rotation, row-major (inverse of the entity's rotation) */
    float translation[3]; /* entity position in world units */
    float reserved;
} fdtd_b200_shape;

/* Rasterise planes [z_begin, z_end) of an nx×ny×nz lattice on `device`: ids_out (host, (z_end-z_begin)*ny*nx u32, x
 * fastest) receives, per cell, the label of the FIRST shape (array order) containing the cell's world point, else
 * default_label.  solver_to_world: 4×4 row-major f64, last row 0 0 0 1 (`transform_from_solver_to_world`). */
int fdtd_b200_rasterize_scene(int device, const uint64_t dims[3], uint64_t z_begin, uint64_t z_end, const double solver_to_world[16],
                              const fdtd_b200_shape* shapes, uint32_t n_shapes, uint32_t default_label, uint32_t* ids_out);
/* The same rasterisation written straight into the handle's device material index (its own slab and, on fused-step
 * slabs, the plane above it) — no host lattice exists at any point.  Labels must be < the handle's table size. */
int fdtd_b200_set_materials_from_scene(fdtd_b200_handle h, const double solver_to_world[16], const fdtd_b200_shape* shapes,
                                       uint32_t n_shapes, uint32_t default_label);

/* ---- field access -------------------------------------------------------------------------- */

/* `Field::field(...).iter()` over the half-open box [lo, hi) in GLOBAL coordinates clipped to the
 * lattice (lib.rs:97-136, util.rs:140-209): out receives (hi-lo) cells × 3 floats, xyz
 * interleaved, x fastest — the fresh field (buffer tick%2, cpu/mod.rs:538).  The box must lie in
 * this handle's slab.  Blocks until the data is on the host. */
int fdtd_b200_read_field(fdtd_b200_handle h, int component, const uint64_t lo[3], const uint64_t hi[3], float* out);
/* The same readback for callers that cannot allocate device memory (code under CUDA graph capture, tight serving loops,
 * harnesses that forbid cudaMalloc).  NO COUNTERPART IN THE REFERENCE.
 * fdtd_b200_read_field gathers the box into a staging buffer it (re)allocates on demand; this variant gathers into a file-scope
 * `static __device__` array of FDTD_B200_STATIC_SCRATCH_BYTES (8 MiB, held by every GPU that loads the library until the process
 * exits) and calls no cudaMalloc / cudaFree / allocating routine.  Same arguments, same result, bit for bit.  Limit: the box may
 * hold at most FDTD_B200_STATIC_SCRATCH_BYTES / 12 = 699 050 cells, else EINVAL (message names the limit) and nothing is read.
 * All calls on a device share the one array, which keeps the previous call's contents (every element that is copied out is
 * overwritten first, so nothing needs clearing): calls that overlap on the GPU — other threads, other handles' streams — race. */
#define FDTD_B200_STATIC_SCRATCH_BYTES 8388608 /* 8 MiB */
int fdtd_b200_read_field_static(fdtd_b200_handle h, int component, const uint64_t lo[3], const uint64_t hi[3], float* out);
/* The inverse (≙ FieldMut, lib.rs:114-127): overwrite a box of E or H. */
int fdtd_b200_write_field(fdtd_b200_handle h, int component, const uint64_t lo[3], const uint64_t hi[3], const float* in);
/* Raw device pointer of one scalar component c in 0..5 = Ex,Ey,Ez,Hx,Hy,Hz (first owned plane).  Cell (x, y, z) of the
 * slab sits at float offset x + pitch·(y + ny·z) with pitch = fdtd_b200_row_pitch(h). */
int fdtd_b200_device_field(fdtd_b200_handle h, int c, void** device_ptr);
/* Cells per stored row.  The reference sizes its lattice as ceil(size/dx) (fdtd/mod.rs:24-29), so nx is rarely a multiple of
 * 16; the float4 / TMA kernels need 16-byte rows of every array, so lattices with nx >= 64 and nx % 16 != 0 are stored with
 * rows padded to the next multiple of 16 (97..127 → 128, the fused kernel's tile; 96 and 112 too, once the lattice is larger than
 * the on-chip kernels' 303 104-cell budget).  The pad cells index an internal all-zero coefficient row and stay
 * 0; they never influence a lattice cell (the x+1 boundary rule is applied at the caller's nx).  Everything that crosses this
 * ABI — linear cell indices of sources and probes, boxes, material ids, PML profiles, projections, the coefficient table —
 * is in terms of the caller's nx × ny × nz lattice; the pitch shows only in fdtd_b200_device_field, in the halo planes
 * (fdtd_b200_halo_plane: pitch·ny floats), in fdtd_b200_memory_required and in the opaque PML state.  Setting the
 * environment variable FDTD_B200_NO_ROW_PADDING=1 disables the padding (pitch == nx). */
uint64_t fdtd_b200_row_pitch(fdtd_b200_handle h);

/* Probe sampling fused into the E half-step: after every step, E and H of each registered cell
 * (global linear index) are appended to a device ring.  Re-registering clears the ring. */
int fdtd_b200_register_probes(fdtd_b200_handle h, uint32_t n, const uint64_t* cells);
/* out[s][k][0..2] = E, [3..5] = H of probe k after step (first_tick + s + 1) was completed, i.e. the
 * sample taken when tick went from first_tick+s to first_tick+s+1.  Probes outside this slab read
 * as zeros.  EINVAL if the range has been overwritten in the ring or is not complete yet. */
int fdtd_b200_read_probes(fdtd_b200_handle h, uint64_t first_tick, uint64_t n_steps, float* out);

/* ---- projection ---------------------------------------------------------------------------- */

enum {
    FDTD_B200_COLORMAP_CPU = 0,    /* cpu/project.rs:117-165: uv=(x,y)/(w+1,h+1), M·(F,1), clamp, sRGB-encode */
    FDTD_B200_COLORMAP_LINEAR = 1, /* project.wgsl:46-75 with the default body: pixel-centre uv, clamp(M·(F,1)) */
    FDTD_B200_COLORMAP_APP_Z = 2   /* cem-app/src/solver/runner.rs:761-780: red/green on value.z */
};
/* `ProjectionPassAdd::add_projection` + `finish` (project.rs:112-133): renders the FRESH field into
 * a w×h RGBA8 image.  projection / color_map: 4×4 column-major (nalgebra storage).  out_rgba is a
 * host buffer of w*h*4 bytes (rgba_is_device = 0) or a device pointer (1).  The z-plane(s) sampled
 * must lie inside this slab; pixels that fall outside it are left untouched. */
int fdtd_b200_project(fdtd_b200_handle h, int component, const float projection[16], const float color_map[16],
                      int colormap_mode, uint32_t w, uint32_t hgt, uint8_t* out_rgba, int rgba_is_device);

#ifdef __cplusplus
}
#endif
#endif /* FDTD_B200_H */
