//! What cem-app gains to run the B200 backend behind its existing solver thread.  Three additions, nothing else in the
//! application changes: the thread keeps calling `begin_update / sources.apply / finish / observers.run`
//! (cem-app/src/solver/runner.rs:471-557).
//!
//! 1. `cem-app/src/solver/config.rs:83-87` — one more variant:
//!
//! ```ignore
//! #[derive(Clone, Copy, Debug, PartialEq, Eq, Serialize, Deserialize)]
//! pub enum Parallelization {
//!     MultiThreaded { num_threads: Option<usize> },
//!     Wgpu,
//!     /// libfdtd_b200.so on this CUDA device (sm_100a)
//!     B200 { device: i32 },
//! }
//! ```
//!
//! 2. `cem-app/src/solver/runner.rs:198-229` (`SolverRunner::run_fdtd`) — one more match arm.  `memory_limit` goes into the
//!    backend so that it can choose the fused step (two field sets) or the in-place step; the runner's own
//!    `memory_required` / `memory_limit` check (runner.rs:304-330) then sees the bytes of the chosen mode:
//!
//! ```ignore
//! Some(Parallelization::B200 { device }) => {
//!     tracing::debug!(device, "using fdtd-b200 (CUDA sm_100a) backend");
//!     let backend = FdtdB200Backend::new(*device).with_memory_limit(common_config.memory_limit);
//!     run_fdtd.run_fdtd_with_backend(&backend)?
//! }
//! ```
//!
//! 3. `cem-app/src/solver/observer.rs` (next to the cpu / wgpu impls at :167-222) — the projection target the runner's
//!    bounds name (`Backend::Instance: CreateProjection<TextureSenderTarget>`, runner.rs:249).  Like the CPU backend the
//!    B200 backend renders into the shared host image of the texture channel (the CUDA kernel's RGBA8 buffer is copied into
//!    it by `fdtd_b200_project`); there is no wgpu/CUDA texture interop.  This file is that addition, compilable inside
//!    cem-app (the orphan rule allows it: `TextureSenderTarget` is local to cem-app).
use cem_solver::project::{CreateProjection, ProjectionParameters, ProjectionPassAdd};
use cem_solver_b200::{
    FdtdB200SolverInstance, FdtdB200SolverState,
    project::{B200ImageProjection, B200ProjectionPass},
};

use crate::solver::observer::{CopyToTextureImageTarget, TextureSenderTarget};

#[derive(Debug)]
pub struct FdtdB200TextureSenderProjection {
    pub projection: B200ImageProjection<CopyToTextureImageTarget>,
}

impl CreateProjection<TextureSenderTarget> for FdtdB200SolverInstance {
    type Projection = FdtdB200TextureSenderProjection;

    fn create_projection(
        &self,
        state: &FdtdB200SolverState,
        target: TextureSenderTarget,
        parameters: &ProjectionParameters,
    ) -> FdtdB200TextureSenderProjection {
        // host image shared with the renderer; marked dirty and uploaded by the render thread
        // (cem-render/src/texture/channel.rs:99-165), exactly as for the CPU backend (observer.rs:167-185)
        let image_sender = target.texture_sender.send_images();
        tracing::debug!(size = ?image_sender.size(), "creating fdtd-b200 projection with image sender");
        let projection = self.create_projection(state, CopyToTextureImageTarget { image_sender }, parameters);
        FdtdB200TextureSenderProjection { projection }
    }
}

impl<'a> ProjectionPassAdd<'a, FdtdB200TextureSenderProjection> for B200ProjectionPass<'a> {
    fn add_projection(&mut self, projection: &'a mut FdtdB200TextureSenderProjection) {
        self.add_projection(&mut projection.projection);
    }
}

// `FdtdB200TextureSenderProjection: Send + 'static` (runner.rs:257-258) holds because `ImageSender` is `Send` (the CPU
// backend's projection, which wraps the same type, already crosses into the solver thread) and `B200ImageProjection` holds
// only host data besides its target.
