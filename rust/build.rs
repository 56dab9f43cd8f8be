// build.rs — links libnerf_b200.so (the reference's build.rs:1-10 only links Accelerate on macOS).
//
// NERF_B200_LIB_DIR: directory holding libnerf_b200.so; default = ../nerf_rs_b200 (where
// `make -C nerf_rs_b200/csrc` puts it: nvcc -gencode arch=compute_100a,code=sm_100a).
use std::env;
use std::path::PathBuf;

fn main() {
    let manifest = PathBuf::from(env::var("CARGO_MANIFEST_DIR").unwrap());
    let dir = env::var("NERF_B200_LIB_DIR")
        .map(PathBuf::from)
        .unwrap_or_else(|_| manifest.join("..").join("nerf_rs_b200"));
    let dir = dir.canonicalize().unwrap_or(dir);
    if !dir.join("libnerf_b200.so").exists() {
        panic!(
            "libnerf_b200.so not found in {} - run `make -C nerf_rs_b200/csrc` or set NERF_B200_LIB_DIR",
            dir.display()
        );
    }
    println!("cargo:rustc-link-search=native={}", dir.display());
    println!("cargo:rustc-link-lib=dylib=nerf_b200");
    println!("cargo:rustc-link-arg=-Wl,-rpath,{}", dir.display());
    println!("cargo:rerun-if-env-changed=NERF_B200_LIB_DIR");
    // the weights the crate bakes in by path (lib.rs:649-650: CARGO_MANIFEST_DIR/lego_rust)
    let data = manifest.join("..").join("data").join("lego_rust");
    println!("cargo:rustc-env=NERF_RS_LEGO_ROOT={}", data.canonicalize().unwrap_or(data).display());
}
