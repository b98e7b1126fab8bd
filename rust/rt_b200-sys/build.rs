// Links librt_b200.so (built by `make -C rust_tracer_b200/csrc`; depends only on libcudart).
// RT_B200_LIB_DIR = the directory that holds it (default: ../../rust_tracer_b200).
use std::env;
use std::path::PathBuf;

fn main() {
    let dir = env::var("RT_B200_LIB_DIR").map(PathBuf::from).unwrap_or_else(|_| {
        PathBuf::from(env::var("CARGO_MANIFEST_DIR").unwrap()).join("../../rust_tracer_b200")
    });
    println!("cargo:rustc-link-search=native={}", dir.display());
    println!("cargo:rustc-link-lib=dylib=rt_b200");
    println!("cargo:rustc-link-arg=-Wl,-rpath,{}", dir.display());
    println!("cargo:rerun-if-env-changed=RT_B200_LIB_DIR");
}
