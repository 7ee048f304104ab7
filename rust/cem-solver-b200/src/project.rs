//! Projections (cem-solver/src/project.rs:75-133) over `fdtd_b200_project`: the CUDA kernel renders the field slice into an
//! RGBA8 device buffer, the library copies it into the target's `image::ImageBuffer` — no graphics API involved.
//!
//! * `CreateProjection<Target>` for every `Target: FdtdImageTarget<Pixel = Rgba<u8>>` (an `RgbaImage`, the reference's
//!   `GifEncoder`, the app's `CopyToTextureImageTarget`), mirroring `FdtdCpuImageProjection` (fdtd/cpu/project.rs:37-75).
//! * `add_projection` renders immediately; target errors and engine errors are collected and `finish` returns them all, as
//!   `FdtdCpuProjectionPass` does (fdtd/cpu/project.rs:168-189).
//! * Colour mapping: `color_map_code = None` → the linear map `clamp(M·(F,1))` of project.wgsl:46-75
//!   (`FDTD_B200_COLORMAP_LINEAR`); the application's `value.z` snippet (cem-app/src/solver/runner.rs:761-780) →
//!   `FDTD_B200_COLORMAP_APP_Z`; any other WGSL snippet is refused with an error (there is no WGSL compiler behind a CUDA
//!   kernel).  [`B200ImageProjection::cpu_style`] selects the CPU backend's variant instead (uv = (x,y)/(w+1,h+1), sRGB
//!   encoding; fdtd/cpu/project.rs:117-165).
//! * The picture shows the field of the CURRENT tick.  The reference reads the swap buffer `tick + 1` in its projection
//!   passes, i.e. the previous step (cpu/project.rs:107, wgpu/project.rs:482); with fields updated in place / ping-pong
//!   committed per step that stale buffer is not part of the state.
use std::os::raw::c_int;

use cem_solver::{
    SolverInstance,
    project::{
        BeginProjectionPass, CreateProjection, FdtdImageTarget, ProjectionParameters, ProjectionPass, ProjectionPassAdd,
    },
};

use crate::{B200Error, FdtdB200SolverInstance, FdtdB200SolverState, error::check_handle, ffi, field::ffi_component};

/// Marker of the application's colour-map snippet (runner.rs:761-780).
const APP_Z_SNIPPET_MARK: &str = "var x = value.z";

#[derive(Debug)]
pub struct B200ImageProjection<Target>
where
    Target: FdtdImageTarget,
{
    target: Target,
    parameters: ProjectionParameters,
    mode: Result<c_int, B200Error>,
}

impl<Target> B200ImageProjection<Target>
where
    Target: FdtdImageTarget,
{
    /// Render like the reference's CPU backend (cpu/project.rs:117-165) instead of like its wgpu backend.
    pub fn cpu_style(mut self) -> Self {
        self.mode = Ok(ffi::FDTD_B200_COLORMAP_CPU);
        self
    }

    pub fn target(&self) -> &Target {
        &self.target
    }

    pub fn into_target(self) -> Target {
        self.target
    }
}

// the projection holds host data and the target only: it is `Send` whenever the target is, which is what the runner
// requires of `Projection` (cem-app/src/solver/runner.rs:257-258)

fn colormap_mode(parameters: &ProjectionParameters) -> Result<c_int, B200Error> {
    match &parameters.color_map_code {
        None => Ok(ffi::FDTD_B200_COLORMAP_LINEAR),
        Some(code) if code.contains(APP_Z_SNIPPET_MARK) => Ok(ffi::FDTD_B200_COLORMAP_APP_Z),
        Some(_) => Err(B200Error::InvalidArgument(
            "color_map_code: only the application's value.z snippet (cem-app/src/solver/runner.rs:761-780) has a CUDA \
             equivalent; pass None for the linear colour map"
                .to_owned(),
        )),
    }
}

impl<Target> CreateProjection<Target> for FdtdB200SolverInstance
where
    Target: FdtdImageTarget<Pixel = image::Rgba<u8>>,
{
    type Projection = B200ImageProjection<Target>;

    fn create_projection(
        &self,
        state: &FdtdB200SolverState,
        target: Target,
        parameters: &ProjectionParameters,
    ) -> B200ImageProjection<Target> {
        let _ = state;
        B200ImageProjection {
            target,
            parameters: parameters.clone(),
            mode: colormap_mode(parameters),
        }
    }
}

impl BeginProjectionPass for FdtdB200SolverInstance {
    type ProjectionPass<'a>
        = B200ProjectionPass<'a>
    where
        Self: 'a;

    fn begin_projection_pass<'a>(&'a self, state: &'a <Self as SolverInstance>::State) -> B200ProjectionPass<'a> {
        B200ProjectionPass {
            state,
            errors: Vec::new(),
        }
    }
}

#[derive(Debug)]
pub struct B200ProjectionPass<'a> {
    state: &'a FdtdB200SolverState,
    errors: Vec<Box<dyn std::error::Error + Send + Sync + 'static>>,
}

impl B200ProjectionPass<'_> {
    fn render(
        &self,
        image: &mut [u8],
        width: u32,
        height: u32,
        parameters: &ProjectionParameters,
        mode: c_int,
    ) -> Result<(), B200Error> {
        if image.len() != 4 * width as usize * height as usize {
            return Err(B200Error::InvalidArgument(format!(
                "image buffer of {} bytes is not {width}x{height} RGBA8",
                image.len()
            )));
        }
        // nalgebra stores matrices column-major, which is what fdtd_b200_project takes
        let rc = unsafe {
            ffi::fdtd_b200_project(
                self.state.h,
                ffi_component(parameters.field),
                parameters.projection.as_slice().as_ptr(),
                parameters.color_map.as_slice().as_ptr(),
                mode,
                width,
                height,
                image.as_mut_ptr(),
                0,
            )
        };
        check_handle(self.state.h, rc)
    }
}

impl<'a, Target> ProjectionPassAdd<'a, B200ImageProjection<Target>> for B200ProjectionPass<'a>
where
    Target: FdtdImageTarget<Pixel = image::Rgba<u8>>,
{
    fn add_projection(&mut self, projection: &'a mut B200ImageProjection<Target>) {
        let mode = match &projection.mode {
            Ok(mode) => *mode,
            Err(error) => {
                self.errors.push(Box::new(error.clone()));
                return;
            }
        };
        let parameters = &projection.parameters;
        let mut engine_error = None;
        let result = projection.target.with_image_buffer(|image| {
            let (width, height) = image.dimensions();
            if let Err(error) = self.render(&mut **image, width, height, parameters, mode) {
                engine_error = Some(error);
            }
        });
        if let Err(error) = result {
            self.errors.push(Box::new(error));
        }
        if let Some(error) = engine_error {
            self.errors.push(Box::new(error));
        }
    }
}

impl ProjectionPass for B200ProjectionPass<'_> {
    type Error = B200ProjectionPassError;

    fn finish(self) -> Result<(), B200ProjectionPassError> {
        // projections are rendered as they are added (the library synchronises on the copy back), nothing is pending
        if self.errors.is_empty() {
            Ok(())
        }
        else {
            Err(B200ProjectionPassError { errors: self.errors })
        }
    }
}

#[derive(Debug, thiserror::Error)]
#[error("fdtd-b200 projection pass error ({} projection(s) failed)", errors.len())]
pub struct B200ProjectionPassError {
    pub errors: Vec<Box<dyn std::error::Error + Send + Sync + 'static>>,
}
