//! Links libfdtd_b200.so.  `FDTD_B200_LIB_DIR` points at the directory that holds it (the `fdtd_b200/` package directory
//! of the fdtd_b200 repository after `python -m fdtd_b200.build`, i.e. `nvcc -gencode arch=compute_100a,code=sm_100a …`).
use std::{env, path::PathBuf};

fn main() {
    println!("cargo:rerun-if-env-changed=FDTD_B200_LIB_DIR");
    let dir = env::var_os("FDTD_B200_LIB_DIR")
        .map(PathBuf::from)
        .unwrap_or_else(|| PathBuf::from("../../fdtd_b200"));
    println!("cargo:rustc-link-search=native={}", dir.display());
    println!("cargo:rustc-link-lib=dylib=fdtd_b200");
    // the library needs only the CUDA runtime, which it links itself (-lcudart); rpath so `cargo run` finds it
    println!("cargo:rustc-link-arg=-Wl,-rpath,{}", dir.display());
}
