// Locates libntcomp_b200.so.  NTCOMP_B200_LIB = the directory that holds it (in this repository: ntcomp_b200/, after
// `make -C ntcomp_b200/csrc`); the run-time search path is baked in so `cargo run` / `cargo test` find it.
use std::env;
use std::path::PathBuf;

fn main() {
    println!("cargo:rerun-if-env-changed=NTCOMP_B200_LIB");
    let dir = env::var("NTCOMP_B200_LIB").map(PathBuf::from).unwrap_or_else(|_| {
        // default: this crate lives in <repo>/rust/ntcomp-b200-sys
        PathBuf::from(env::var("CARGO_MANIFEST_DIR").unwrap()).join("../../ntcomp_b200")
    });
    let dir = dir.canonicalize().unwrap_or_else(|_| panic!("NTCOMP_B200_LIB: no such directory: {}", dir.display()));
    assert!(dir.join("libntcomp_b200.so").exists(), "libntcomp_b200.so not found in {} (run `make -C ntcomp_b200/csrc`)", dir.display());
    println!("cargo:rustc-link-search=native={}", dir.display());
    println!("cargo:rustc-link-lib=dylib=ntcomp_b200");
    println!("cargo:rustc-link-arg=-Wl,-rpath,{}", dir.display());
}
