//! Status codes of the C ABI as a Rust error.  The reference's backends use `Infallible` for `create_instance`
//! (cem-solver/src/fdtd/cpu/mod.rs:200, wgpu/mod.rs:112); a CUDA backend can fail (no device, out of memory), so this one
//! has a real error type.  `set_forcing` on an out-of-lattice point panics exactly like the reference
//! (cpu/mod.rs:392-396, wgpu/mod.rs:342-346) — that is done in lib.rs, not here.
use std::{ffi::CStr, os::raw::c_int};

use crate::ffi;

#[derive(Clone, Debug, PartialEq, Eq, thiserror::Error)]
pub enum B200Error {
    #[error("fdtd_b200: invalid argument: {0}")]
    InvalidArgument(String),
    #[error("fdtd_b200: out of device memory: {0}")]
    OutOfMemory(String),
    #[error("fdtd_b200: CUDA error or no usable device (there is no CPU fallback): {0}")]
    Cuda(String),
    #[error("fdtd_b200: call not valid in the handle's current state: {0}")]
    State(String),
    #[error("fdtd_b200: the PML described by DomainDescription::pml is not separable: {0}")]
    PmlNotSeparable(String),
    #[error("fdtd_b200: library ABI version {found}, this crate was written for {expected}")]
    AbiMismatch { found: c_int, expected: c_int },
    #[error("fdtd_b200: unknown status code {0}: {1}")]
    Unknown(c_int, String),
}

fn c_message(p: *const std::os::raw::c_char) -> String {
    if p.is_null() {
        return String::new();
    }
    // SAFETY: the library returns a NUL-terminated string that stays valid until the next failing call on this thread /
    // handle; it is copied out immediately.
    unsafe { CStr::from_ptr(p) }.to_string_lossy().into_owned()
}

/// Message of the last failing call made without a handle (`fdtd_b200_global_error`).
pub fn global_error() -> String {
    c_message(unsafe { ffi::fdtd_b200_global_error() })
}

/// Message of the last failing call on `h` (`fdtd_b200_last_error`).
pub fn last_error(h: ffi::fdtd_b200_handle) -> String {
    if h.is_null() {
        return global_error();
    }
    c_message(unsafe { ffi::fdtd_b200_last_error(h) })
}

fn from_code(rc: c_int, message: String) -> B200Error {
    match rc {
        ffi::FDTD_B200_EINVAL => B200Error::InvalidArgument(message),
        ffi::FDTD_B200_ENOMEM => B200Error::OutOfMemory(message),
        ffi::FDTD_B200_ECUDA => B200Error::Cuda(message),
        ffi::FDTD_B200_ESTATE => B200Error::State(message),
        other => B200Error::Unknown(other, message),
    }
}

/// `rc` of a handle-less call → `Result`.
pub fn check(rc: c_int) -> Result<(), B200Error> {
    if rc == ffi::FDTD_B200_OK {
        Ok(())
    }
    else {
        Err(from_code(rc, global_error()))
    }
}

/// `rc` of a call on handle `h` → `Result`.
pub fn check_handle(h: ffi::fdtd_b200_handle, rc: c_int) -> Result<(), B200Error> {
    if rc == ffi::FDTD_B200_OK {
        Ok(())
    }
    else {
        Err(from_code(rc, last_error(h)))
    }
}

/// Refuses a library of another ABI generation at first use.
pub fn check_abi() -> Result<(), B200Error> {
    let found = unsafe { ffi::fdtd_b200_abi_version() };
    if found == ffi::FDTD_B200_ABI_VERSION {
        Ok(())
    }
    else {
        Err(B200Error::AbiMismatch {
            found,
            expected: ffi::FDTD_B200_ABI_VERSION,
        })
    }
}
